// cepstrum_kernel.h -- launch interface of the fused cepstrum (MFCC stats + SVM) kernel.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "sj_tables.h"

struct SjCepstrumArgs {
    const float *clips;   // [n_clips][n_samples], device
    long n_clips;
    int n_samples;
    const SjTables *tables; // device copy
    int32_t *labels;      // [n_clips] or NULL
    float *conf;          // [n_clips] or NULL
    float *score;         // [n_clips] or NULL
    float *feats;         // [n_clips][40] or NULL
    int do_svm;           // 0: features only
};

// Enqueues the kernel on `stream`.  `host_tables` is the host copy of a.tables (selects the mel layout),
// `num_sms` the SM count of the device.  Returns the cudaError_t of the launch.
cudaError_t sj_launch_cepstrum(const SjCepstrumArgs &a, const SjTables *host_tables, int num_sms, cudaStream_t stream);
// one-time attribute set-up (max dynamic shared memory); returns cudaError_t
cudaError_t sj_cepstrum_setup(int mel_maxseg);
// batched standalone SVM: feats [n][40] -> labels/conf/score
cudaError_t sj_launch_svm(const float *feats, long n, const SjTables *tables, int32_t *labels, float *conf,
                          float *score, cudaStream_t stream);
