// sampler.cu — on-device token selection + bookkeeping of the
// device-resident decode state, so that a whole decode loop replays as one CUDA graph with no
// host round trip per token.
//
// Chain (llama-server's default order, of which the reference only sets `temperature` and
// `max_tokens`, runtime/src/inference.rs:114-122): top-k -> top-p -> min-p -> temperature ->
// softmax -> draw.  temperature <= 0 selects greedy argmax with the lowest index winning ties
// (oracle/aios_oracle.c argmax_first).
#include "common.cuh"
#include "kernels.h"

namespace aios {

constexpr int SAMP_THREADS = 1024;
constexpr int SAMP_MAX_K = 128;

__device__ __forceinline__ unsigned long long pack_key(float v, int idx) {
    uint32_t b = __float_as_uint(v);
    b = (b & 0x80000000u) ? ~b : (b | 0x80000000u);
    return ((unsigned long long)b << 32) | (uint32_t)(0x7fffffff - idx);
}
__device__ __forceinline__ float key_value(unsigned long long k) {
    uint32_t b = (uint32_t)(k >> 32);
    b = (b & 0x80000000u) ? (b & 0x7fffffffu) : ~b;
    return __uint_as_float(b);
}
__device__ __forceinline__ int key_index(unsigned long long k) { return 0x7fffffff - (int)(uint32_t)(k & 0xffffffffu); }

__device__ __forceinline__ unsigned long long block_max_key(unsigned long long k, unsigned long long *sh) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long t = __shfl_xor_sync(0xffffffffu, k, o);
        k = t > k ? t : k;
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __syncthreads();
    if (lane == 0) sh[warp] = k;
    __syncthreads();
    if (warp == 0) {
        k = sh[lane];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const unsigned long long t = __shfl_xor_sync(0xffffffffu, k, o);
            k = t > k ? t : k;
        }
        if (lane == 0) sh[0] = k;
    }
    __syncthreads();
    return sh[0];
}

// splitmix64 -> uniform in [0,1)
__device__ __forceinline__ float rng_uniform(uint64_t seed, uint64_t ctr) {
    uint64_t z = seed + 0x9E3779B97F4A7C15ull * (ctr + 1);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z ^= z >> 31;
    return (float)(z >> 40) * (1.0f / 16777216.0f);
}

__global__ void __launch_bounds__(SAMP_THREADS) sample_kernel(const SampleParams p) {
    __shared__ unsigned long long sh[32];
    __shared__ unsigned long long top[SAMP_MAX_K];
    __shared__ int chosen;
    const int V = p.n_vocab;
    const bool greedy = !(p.temperature > 0.f);
    const int k = greedy ? 1 : (p.top_k > 0 ? min(p.top_k, SAMP_MAX_K) : SAMP_MAX_K);
    // k rounds of block arg-max over the candidates not yet taken (keys are unique per index)
    unsigned long long bound = ~0ull;
    for (int r = 0; r < k; r++) {
        unsigned long long best = 0ull;
        for (int i = threadIdx.x; i < V; i += SAMP_THREADS) {
            const unsigned long long key = pack_key(p.logits[i], i);
            if (key < bound && key > best) best = key;
        }
        best = block_max_key(best, sh);
        if (threadIdx.x == 0) top[r] = best;
        bound = best;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        int tok = key_index(top[0]);
        if (!greedy) {
            // the random draw is a function of (seed, absolute position): the same whether the caller decodes in
            // one call or in chunks
            const int step = p.pos_ptr ? *p.pos_ptr : *p.step_ptr;
            // softmax over the k candidates (sorted descending), then top-p / min-p cut, then temperature
            const float mx = key_value(top[0]);
            float probs[SAMP_MAX_K];
            float sum = 0.f;
            for (int i = 0; i < k; i++) { probs[i] = expf(key_value(top[i]) - mx); sum += probs[i]; }
            int n = k;
            if (p.top_p < 1.f) {
                float c = 0.f;
                for (int i = 0; i < k; i++) { c += probs[i] / sum; if (c >= p.top_p) { n = i + 1; break; } }
            }
            if (p.min_p > 0.f) {
                int n2 = 1;
                while (n2 < n && probs[n2] >= p.min_p * probs[0]) n2++;
                n = n2;
            }
            float sum2 = 0.f;
            for (int i = 0; i < n; i++) { probs[i] = expf((key_value(top[i]) - mx) / p.temperature); sum2 += probs[i]; }
            const float u = rng_uniform(p.seed, (uint64_t)step) * sum2;
            float c = 0.f;
            int pick = n - 1;
            for (int i = 0; i < n; i++) { c += probs[i]; if (u < c) { pick = i; break; } }
            tok = key_index(top[pick]);
        }
        chosen = tok;
        const int step = *p.step_ptr;
        if (step < p.out_cap) p.out_tokens[step] = tok;
        *p.step_ptr = step + 1;
        *p.cur_token = tok;
        if (p.pos_ptr) *p.pos_ptr += 1;  // the sampled token is fed at the next position
    }
    __syncthreads();
    (void)chosen;
}

void launch_sample(const SampleParams &p, cudaStream_t st) { sample_kernel<<<1, SAMP_THREADS, 0, st>>>(p); }

}  // namespace aios
