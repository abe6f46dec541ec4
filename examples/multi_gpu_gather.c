/*
 * multi_gpu_gather.c -- the multi-GPU driver of libscrubjay_cuda in plain C: one PROCESS per GPU, clips sharded by
 * contiguous ranges, the label / score gather fused into the kernel as NVLink peer stores.  No torch, no
 * NCCL, no MPI: the 64-byte CUDA IPC handles travel over pipes to the parent and back.
 *
 *   usage: multi_gpu_gather [n_gpus (default 2)] [clips per GPU (default 4096)]
 *
 * Protocol (each child r of W):
 *   1. scrubjay_create on device r; scrubjay_peer_alloc two result arrays for ALL W * n clips (labels int32, score fp32)
 *   2. send both handles to the parent; receive the handles of all ranks
 *   3. scrubjay_peer_open the other ranks' arrays; scrubjay_gather_bind(all W targets, out_base = r * n)
 *   4. barrier (everybody bound); scrubjay_infer_batch_gather_dev on the own shard; wait for the kernel; barrier
 *   5. check: every shard slot of the own gathered arrays against a checksum each owner computed from a plain
 *      scrubjay_infer_batch_dev run of its shard
 * Build: examples/Makefile (gcc + libcudart for the two host copies of the check; the library itself is plain C ABI).
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/types.h>
#include <sys/wait.h>
#include <unistd.h>

#include <cuda_runtime_api.h>

#include "../include/scrubjay_cuda.h"

#define MAX_RANKS SJ_MAX_PEERS
#define N_SAMPLES 16000

typedef struct { unsigned char labels[SJ_PEER_HANDLE_BYTES], score[SJ_PEER_HANDLE_BYTES]; } handle_pair;
typedef struct { uint64_t label_sum, score_hash; } shard_sum;

static int read_all(int fd, void *p, size_t n)
{
    char *c = (char *)p;
    while (n) { ssize_t r = read(fd, c, n); if (r <= 0) return -1; c += r; n -= (size_t)r; }
    return 0;
}
static int write_all(int fd, const void *p, size_t n)
{
    const char *c = (const char *)p;
    while (n) { ssize_t r = write(fd, c, n); if (r <= 0) return -1; c += r; n -= (size_t)r; }
    return 0;
}

/* order-independent checksum of one shard of the result arrays (host copies) */
static shard_sum checksum(const int32_t *labels, const float *score, long n)
{
    shard_sum s = {0, 0};
    for (long i = 0; i < n; i++) {
        uint32_t bits;
        memcpy(&bits, &score[i], 4);
        s.label_sum += (uint64_t)labels[i];
        s.score_hash += (uint64_t)bits * 0x9E3779B97F4A7C15ull + (uint64_t)(i + 1) * bits;
    }
    return s;
}

#define CHECK(call)                                                                                           \
    do {                                                                                                      \
        int rc_ = (call);                                                                                     \
        if (rc_ != SJ_OK) { fprintf(stderr, "rank %d: %s failed (%d): %s\n", rank, #call, rc_, scrubjay_last_error(ctx)); return 1; } \
    } while (0)

/* all-to-all of fixed-size items through the parent: send mine, receive everybody's (also a barrier) */
static int exchange(int to_parent, int from_parent, const void *mine, void *all, size_t item, int world)
{
    if (write_all(to_parent, mine, item)) return -1;
    return read_all(from_parent, all, item * (size_t)world);
}

static int child(int rank, int world, long n, int to_parent, int from_parent)
{
    scrubjay_ctx *ctx = NULL;
    int rc = scrubjay_create(&ctx, rank, NULL, 0, 16000);
    if (rc != SJ_OK) { fprintf(stderr, "rank %d: scrubjay_create failed (%d): %s\n", rank, rc, scrubjay_last_error(NULL)); return 1; }
    const long total = n * world;
    void *d_labels = NULL, *d_score = NULL, *d_clips = NULL, *d_own_labels = NULL, *d_own_score = NULL;
    handle_pair mine, all[MAX_RANKS];
    CHECK(scrubjay_peer_alloc(ctx, (size_t)total * 4, &d_labels, mine.labels));
    CHECK(scrubjay_peer_alloc(ctx, (size_t)total * 4, &d_score, mine.score));
    /* plain device allocations through the same call (no handle wanted) */
    CHECK(scrubjay_peer_alloc(ctx, (size_t)n * N_SAMPLES * 4, &d_clips, NULL));
    CHECK(scrubjay_peer_alloc(ctx, (size_t)n * 4, &d_own_labels, NULL));
    CHECK(scrubjay_peer_alloc(ctx, (size_t)n * 4, &d_own_score, NULL));
    cudaMemset(d_labels, 0xff, (size_t)total * 4); /* poison: a lost peer store shows up in the checksums */
    cudaMemset(d_score, 0xff, (size_t)total * 4);
    cudaDeviceSynchronize();
    if (exchange(to_parent, from_parent, &mine, all, sizeof mine, world)) return 1;

    int32_t *pl[MAX_RANKS];
    float *ps[MAX_RANKS];
    for (int r = 0; r < world; r++) {
        if (r == rank) { pl[r] = (int32_t *)d_labels; ps[r] = (float *)d_score; continue; }
        void *a = NULL, *b = NULL;
        CHECK(scrubjay_peer_open(ctx, all[r].labels, &a));
        CHECK(scrubjay_peer_open(ctx, all[r].score, &b));
        pl[r] = (int32_t *)a; ps[r] = (float *)b;
    }
    CHECK(scrubjay_gather_bind(ctx, world, pl, ps, (long)rank * n));

    /* the own shard: clips [rank * n, (rank + 1) * n) of the counter-based synthetic set */
    CHECK(scrubjay_synth_clips_dev(ctx, (float *)d_clips, (long)rank * n, n, N_SAMPLES, 0x5C2B));
    CHECK(scrubjay_infer_batch_dev(ctx, (const float *)d_clips, n, N_SAMPLES, (int32_t *)d_own_labels, NULL, (float *)d_own_score, NULL));
    char token = 1, tokens[MAX_RANKS];
    if (exchange(to_parent, from_parent, &token, tokens, 1, world)) return 1; /* barrier: every rank has bound its targets */
    CHECK(scrubjay_infer_batch_gather_dev(ctx, (const float *)d_clips, n, N_SAMPLES, NULL, NULL));
    float ms = 0.f;
    CHECK(scrubjay_last_kernel_ms(ctx, &ms)); /* waits for the kernel: its peer stores have drained */
    if (exchange(to_parent, from_parent, &token, tokens, 1, world)) return 1; /* barrier: the kernels of ALL ranks are done */

    int32_t *h_labels = (int32_t *)malloc((size_t)total * 4), *h_own_l = (int32_t *)malloc((size_t)n * 4);
    float *h_score = (float *)malloc((size_t)total * 4), *h_own_s = (float *)malloc((size_t)n * 4);
    cudaMemcpy(h_labels, d_labels, (size_t)total * 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(h_score, d_score, (size_t)total * 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(h_own_l, d_own_labels, (size_t)n * 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(h_own_s, d_own_score, (size_t)n * 4, cudaMemcpyDeviceToHost);
    shard_sum own = checksum(h_own_l, h_own_s, n), want[MAX_RANKS];
    if (exchange(to_parent, from_parent, &own, want, sizeof own, world)) return 1;
    int bad = 0;
    for (int r = 0; r < world; r++) {
        shard_sum got = checksum(h_labels + (long)r * n, h_score + (long)r * n, n);
        if (got.label_sum != want[r].label_sum || got.score_hash != want[r].score_hash) {
            fprintf(stderr, "rank %d: shard %d of the gathered arrays differs from rank %d's plain run\n", rank, r, r);
            bad = 1;
        }
    }
    printf("rank %d: device %d, %ld clips, fused-gather kernel %.3f ms, %d shards %s\n", rank, rank, n, ms, world,
           bad ? "MISMATCH" : "verified");
    fflush(stdout);

    /* close protocol: nobody frees while a peer could still write (all kernels are done: barrier above) */
    CHECK(scrubjay_gather_bind(ctx, 0, NULL, NULL, 0));
    for (int r = 0; r < world; r++)
        if (r != rank) { CHECK(scrubjay_peer_close(ctx, pl[r])); CHECK(scrubjay_peer_close(ctx, ps[r])); }
    if (exchange(to_parent, from_parent, &token, tokens, 1, world)) return 1; /* nobody maps these arrays any more */
    scrubjay_peer_free(ctx, d_labels); scrubjay_peer_free(ctx, d_score); scrubjay_peer_free(ctx, d_clips);
    scrubjay_peer_free(ctx, d_own_labels); scrubjay_peer_free(ctx, d_own_score);
    free(h_labels); free(h_score); free(h_own_l); free(h_own_s);
    scrubjay_destroy(ctx);
    return bad;
}

int main(int argc, char **argv)
{
    int world = argc > 1 ? atoi(argv[1]) : 2;
    long n = argc > 2 ? atol(argv[2]) : 4096;
    if (world < 1 || world > MAX_RANKS || n < 1) { fprintf(stderr, "usage: %s [n_gpus 1..8] [clips per GPU]\n", argv[0]); return 2; }
    int up[MAX_RANKS][2], down[MAX_RANKS][2];
    pid_t pid[MAX_RANKS];
    /* fork BEFORE any CUDA call: a CUDA context does not survive fork() */
    for (int r = 0; r < world; r++) {
        if (pipe(up[r]) || pipe(down[r])) { perror("pipe"); return 2; }
        pid[r] = fork();
        if (pid[r] < 0) { perror("fork"); return 2; }
        if (pid[r] == 0) {
            /* the parent's ends of all pipes made so far (the parent has already closed its copies of the children's ends) */
            for (int q = 0; q <= r; q++) { close(up[q][0]); close(down[q][1]); }
            _exit(child(r, world, n, up[r][1], down[r][0]));
        }
        close(up[r][1]); close(down[r][0]);
    }
    /* the parent is the switchboard: rounds of (collect one item per rank, send all items to every rank).
     * Item sizes of the rounds, in the order the children call exchange(). */
    const size_t rounds[] = {sizeof(handle_pair), 1, 1, sizeof(shard_sum), 1};
    int failed = 0;
    for (size_t k = 0; k < sizeof rounds / sizeof rounds[0] && !failed; k++) {
        unsigned char buf[MAX_RANKS * sizeof(handle_pair)];
        for (int r = 0; r < world; r++)
            if (read_all(up[r][0], buf + (size_t)r * rounds[k], rounds[k])) { failed = 1; break; }
        for (int r = 0; r < world && !failed; r++)
            if (write_all(down[r][1], buf, rounds[k] * (size_t)world)) failed = 1;
    }
    for (int r = 0; r < world; r++) { close(up[r][0]); close(down[r][1]); }
    int rc = failed;
    for (int r = 0; r < world; r++) {
        int st = 0;
        waitpid(pid[r], &st, 0);
        if (!WIFEXITED(st) || WEXITSTATUS(st) != 0) rc = 1;
    }
    printf("%s: %d GPU(s) x %ld clips, fused peer-store gather, handles exchanged over pipes\n", rc ? "FAILED" : "OK", world, n);
    return rc;
}
