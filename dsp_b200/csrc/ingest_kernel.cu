// ingest_kernel.cu -- the decode stage in front of the hot path: interleaved int16 PCM
// frames -> mono float in [-1, 1).  Two channel policies, both from the reference:
//   SJ_MIX_FIRST  channel 0 only, x = s / 32768   (read_wav_file + normalisation, donut-classifier/classifier.c:
//                 293-297, 56-60)
//   SJ_MIX_MEAN   mean of the channels of s / 32768 (aubio_source_do's down-mix, used by cepstrum/scrubjay_infer.c:41)
// Pure streaming: 2*channels bytes in, 4 bytes out per frame; grid-stride over frames, coalesced 16-bit loads.
// Also the two firmware-side conversions that feed classify(): 12-bit ADC counts -> float (sync/sync.cpp:236-241)
// and the linear up-sampler (microphone/src/microphone.cpp:55-70).
#include <cuda_runtime.h>
#include <stdint.h>

#include "ingest_kernel.h"

namespace {

__global__ void __launch_bounds__(256) decode_pcm16_kernel(const int16_t *in, long frames, int channels, int policy, float *out)
{
    const long stride = (long)gridDim.x * blockDim.x;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < frames; i += stride) {
        const int16_t *f = in + i * channels;
        float x;
        if (policy == 0 || channels == 1) x = (float)f[0] * (1.f / 32768.f);
        else {
            // aubio: each channel converted to float first, summed in channel order, divided by the count
            float acc = 0.f;
            for (int c = 0; c < channels; ++c) acc = __fadd_rn(acc, (float)f[c] * (1.f / 32768.f));
            x = __fdiv_rn(acc, (float)channels);
        }
        out[i] = x;
    }
}

// 12-bit ADC counts -> float, the firmware's two steps (sync/sync.cpp:236-241): int16 pcm = (adc - 2048) * 16, then
// pcm / 32768.0 stored to a float (exact: |pcm| < 2^15 and the divisor is a power of two)
__global__ void __launch_bounds__(256) decode_adc12_kernel(const uint16_t *in, long n, float *out)
{
    const long stride = (long)gridDim.x * blockDim.x;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const int16_t pcm = (int16_t)(((int)in[i] - 2048) * 16);
        out[i] = (float)((double)pcm / 32768.0);
    }
}

// upsampleLinear (microphone/src/microphone.cpp:55-70) for a batch of buffers: every output sample is independent;
// float arithmetic in the reference's order, multiply and add rounded separately
__global__ void __launch_bounds__(256) upsample_linear_kernel(const float *in, long n_signals, int old_size, int new_size,
                                                              float *out)
{
    const long total = n_signals * (long)new_size, stride = (long)gridDim.x * blockDim.x;
    const float ratio = __fdiv_rn((float)(old_size - 1), (float)(new_size - 1));
    for (long g = (long)blockIdx.x * blockDim.x + threadIdx.x; g < total; g += stride) {
        const long s = g / new_size;
        const int i = (int)(g - s * new_size);
        const float old_index = __fmul_rn((float)i, ratio);
        const int lo = (int)floor((double)old_index);
        const int hi = lo == old_size - 1 ? old_size - 1 : lo + 1;
        const float frac = __fsub_rn(old_index, (float)lo);
        const float *x = in + s * (long)old_size;
        const float v0 = __ldg(x + lo), v1 = __ldg(x + hi);
        out[g] = __fadd_rn(v0, __fmul_rn(__fsub_rn(v1, v0), frac));
    }
}

} // namespace

cudaError_t sj_launch_decode_adc12(const uint16_t *in, long n, float *out, int num_sms, cudaStream_t stream)
{
    if (n <= 0) return cudaSuccess;
    long blocks = (n + 255) / 256;
    if (blocks > (long)num_sms * 8) blocks = (long)num_sms * 8;
    decode_adc12_kernel<<<(unsigned)blocks, 256, 0, stream>>>(in, n, out);
    return cudaGetLastError();
}

cudaError_t sj_launch_upsample_linear(const float *in, long n_signals, int old_size, int new_size, float *out, int num_sms,
                                      cudaStream_t stream)
{
    const long total = n_signals * (long)new_size;
    if (total <= 0) return cudaSuccess;
    long blocks = (total + 255) / 256;
    if (blocks > (long)num_sms * 8) blocks = (long)num_sms * 8;
    upsample_linear_kernel<<<(unsigned)blocks, 256, 0, stream>>>(in, n_signals, old_size, new_size, out);
    return cudaGetLastError();
}

cudaError_t sj_launch_decode_pcm16(const int16_t *in, long frames, int channels, int policy, float *out, int num_sms,
                                   cudaStream_t stream)
{
    if (frames <= 0) return cudaSuccess;
    long blocks = (frames + 255) / 256;
    if (blocks > (long)num_sms * 8) blocks = (long)num_sms * 8;
    decode_pcm16_kernel<<<(unsigned)blocks, 256, 0, stream>>>(in, frames, channels, policy, out);
    return cudaGetLastError();
}
