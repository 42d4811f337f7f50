// nm_fasta.cuh -- what nm_lib.cu needs from nm_fasta.cu (not part of the C-ABI)
#pragma once
#include <stdint.h>

struct nm_fasta;
// device pointer to the parsed bases (all records back to back), their number, and the device they live on
const uint8_t *nm_fasta_device_sequence(const nm_fasta *fa, uint64_t *length, int *device);
// records a CUDA failure where nm_last_cuda_error() finds it (defined in nm_lib.cu)
void nm_internal_set_cuda_error(const char *what, const char *message);
