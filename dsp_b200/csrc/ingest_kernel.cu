// ingest_kernel.cu -- the decode stage in front of the hot path: interleaved int16 PCM
// frames -> mono float in [-1, 1).  Two channel policies, both from the reference:
//   SJ_MIX_FIRST  channel 0 only, x = s / 32768   (read_wav_file + normalisation, donut-classifier/classifier.c:
//                 293-297, 56-60)
//   SJ_MIX_MEAN   mean of the channels of s / 32768 (aubio_source_do's down-mix, used by cepstrum/scrubjay_infer.c:41)
// Pure streaming: 2*channels bytes in, 4 bytes out per frame; grid-stride over frames, coalesced 16-bit loads.
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

} // namespace

cudaError_t sj_launch_decode_pcm16(const int16_t *in, long frames, int channels, int policy, float *out, int num_sms,
                                   cudaStream_t stream)
{
    if (frames <= 0) return cudaSuccess;
    long blocks = (frames + 255) / 256;
    if (blocks > (long)num_sms * 8) blocks = (long)num_sms * 8;
    decode_pcm16_kernel<<<(unsigned)blocks, 256, 0, stream>>>(in, frames, channels, policy, out);
    return cudaGetLastError();
}
