// ingest_kernel.h -- launch interface of the PCM decode stage
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

cudaError_t sj_launch_decode_pcm16(const int16_t *in, long frames, int channels, int policy, float *out, int num_sms,
                                   cudaStream_t stream);
cudaError_t sj_launch_decode_adc12(const uint16_t *in, long n, float *out, int num_sms, cudaStream_t stream);
cudaError_t sj_launch_upsample_linear(const float *in, long n_signals, int old_size, int new_size, float *out, int num_sms,
                                      cudaStream_t stream);
