// ingest_kernel.h -- launch interface of the PCM decode stage
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

cudaError_t sj_launch_decode_pcm16(const int16_t *in, long frames, int channels, int policy, float *out, int num_sms,
                                   cudaStream_t stream);
