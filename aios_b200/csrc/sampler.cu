// sampler.cu — on-device token selection + bookkeeping of the
// device-resident decode state, so that a whole decode loop replays as one CUDA graph with no
// host round trip per token.
//
// Chain (llama-server's default order, of which the reference only sets `temperature` and
// `max_tokens`, runtime/src/inference.rs:114-122): [logit mask] -> top-k -> top-p -> min-p ->
// temperature -> softmax -> draw.  temperature <= 0 selects greedy argmax with the lowest index
// winning ties.  The arithmetic is oracle/aios_oracle.c oracle_sample(), operation by operation
// (single rounded f32 steps, the deterministic expf_det, the same splitmix64 uniform): candidate
// set, both cut-offs and the drawn token are bit-identical to the oracle's.
//
// Selection of the K best candidates is a radix select on the packed 64-bit (value, ~index) keys —
// 8-bit digits from the top, one shared-memory histogram per digit, early exit as soon as the bin
// that holds the K-th key is wholly needed (3-4 sweeps of the L2-resident logits in practice) —
// instead of round 1's K full arg-max sweeps (40 x ~2.5 us per token).
//
// The optional mask (one bit per token id; JSON mode: the grammar's allowed set for this step)
// removes tokens before anything else, like llama.cpp's grammar sampler.
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

// splitmix64 -> uniform in [0,1)  (oracle_uniform)
__device__ __forceinline__ float rng_uniform(uint64_t seed, uint64_t ctr) {
    uint64_t z = seed + 0x9E3779B97F4A7C15ull * (ctr + 1);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z ^= z >> 31;
    return __fmul_rn((float)(z >> 40), 1.0f / 16777216.0f);
}

__device__ __forceinline__ bool allowed(const uint8_t *mask, int i) { return !mask || ((mask[i >> 3] >> (i & 7)) & 1); }

__device__ __forceinline__ void sample_body(const SampleParams &p) {
    __shared__ unsigned long long sh[32];
    __shared__ unsigned long long top[SAMP_MAX_K], sorted[SAMP_MAX_K];
    __shared__ float prob[SAMP_MAX_K], wgt[SAMP_MAX_K];
    __shared__ unsigned hist[256];
    __shared__ unsigned long long sel_prefix;
    __shared__ int sel_shift, sel_need, sel_done, n_top;
    const int V = p.n_vocab;
    const bool greedy = !(p.temperature > 0.f);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    int K = 1;

    if (greedy) {
        unsigned long long best = 0ull;
        for (int i = tid; i < V; i += SAMP_THREADS)
            if (allowed(p.mask, i)) {
                const unsigned long long key = pack_key(p.logits[i], i);
                best = key > best ? key : best;
            }
        best = block_max_key(best, sh);
        if (tid == 0) { sorted[0] = best; n_top = best ? 1 : 0; }
        __syncthreads();
        K = n_top;  // 0: the mask allows nothing
    } else {
        K = p.top_k > 0 ? min(p.top_k, SAMP_MAX_K) : SAMP_MAX_K;
        // ---- radix select: T = the K-th largest key; afterwards every key >= T is a candidate
        if (tid == 0) { sel_prefix = 0ull; sel_shift = 56; sel_need = K; sel_done = 0; n_top = 0; }
        __syncthreads();
        for (int pass = 0; pass < 8; pass++) {
            if (tid < 256) hist[tid] = 0u;
            __syncthreads();
            const int shift = sel_shift;
            const unsigned long long prefix = sel_prefix;
            for (int i = tid; i < V; i += SAMP_THREADS)
                if (allowed(p.mask, i)) {
                    const unsigned long long key = pack_key(p.logits[i], i);
                    if (pass == 0 || (key >> (shift + 8)) == prefix) atomicAdd(&hist[(unsigned)(key >> shift) & 255u], 1u);
                }
            __syncthreads();
            if (warp == 0) {
                // suffix sums over the 256 bins, from the top: lane l owns bins [8l, 8l+8)
                unsigned c[8], mine = 0;
#pragma unroll
                for (int j = 0; j < 8; j++) { c[j] = hist[lane * 8 + j]; mine += c[j]; }
                unsigned suf = mine;  // inclusive suffix scan over the lanes
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const unsigned t = __shfl_down_sync(0xffffffffu, suf, o);
                    if (lane + o < 32) suf += t;
                }
                const unsigned above = suf - mine;  // keys in bins of higher lanes
                int need = sel_need;
                const unsigned total = __shfl_sync(0xffffffffu, suf, 0);
                if (pass == 0 && (unsigned)need > total) need = (int)total;  // fewer allowed tokens than K
                // the lane whose bins contain the need-th key from the top
                const bool here = need > 0 && above < (unsigned)need && above + mine >= (unsigned)need;
                const unsigned who = __ballot_sync(0xffffffffu, here);
                if (who == 0u) {
                    if (lane == 0) { sel_done = 1; sel_need = 0; }  // nothing allowed at all
                } else if (lane == __ffs(who) - 1) {
                    unsigned acc = above;
                    int b = 7;
                    for (; b >= 0; b--) {
                        if (acc + c[b] >= (unsigned)need) break;
                        acc += c[b];
                    }
                    const int rem = need - (int)acc;  // keys still to take from bin b
                    sel_prefix = (prefix << 8) | (unsigned long long)(lane * 8 + b);
                    if (pass == 0) n_top = need;
                    if ((unsigned)rem == c[b] || shift == 0) sel_done = 1;  // the whole bin is needed: threshold found
                    else { sel_need = rem; }
                    sel_shift = shift - 8;
                }
            }
            __syncthreads();
            if (sel_done) break;
        }
        // keys >= T, T = sel_prefix << (sel_shift + 8)
        K = n_top;
        if (K > 0) {
            const int sh_bits = sel_shift + 8;
            const unsigned long long T = sel_prefix << sh_bits;
            __syncthreads();
            if (tid == 0) n_top = 0;
            __syncthreads();
            for (int i = tid; i < V; i += SAMP_THREADS)
                if (allowed(p.mask, i)) {
                    const unsigned long long key = pack_key(p.logits[i], i);
                    if (key >= T) {
                        const int slot = atomicAdd(&n_top, 1);
                        if (slot < SAMP_MAX_K) top[slot] = key;
                    }
                }
            __syncthreads();
            K = min(n_top, SAMP_MAX_K);
            // rank sort (keys are unique): descending
            if (tid < K) {
                const unsigned long long mine = top[tid];
                int rank = 0;
                for (int j = 0; j < K; j++) rank += top[j] > mine ? 1 : 0;
                sorted[rank] = mine;
            }
            __syncthreads();
        }
    }

    // ---- the chain on the K sorted candidates (oracle_sample, same operations in the same order)
    if (!greedy && K > 0 && tid < K) {
        const float l0 = key_value(sorted[0]);
        const float d = __fsub_rn(key_value(sorted[tid]), l0);
        prob[tid] = expf_det(d);
        wgt[tid] = expf_det(__fdiv_rn(d, p.temperature));
    }
    __syncthreads();
    if (tid == 0) {
        int tok = K > 0 ? key_index(sorted[0]) : 0;
        int n = K, n_p = K, n_m = K;
        if (!greedy && K > 0) {
            // the random draw is a function of (seed, absolute position): the same whether the caller decodes in
            // one call or in chunks
            const int ctr = p.pos_ptr ? *p.pos_ptr : *p.step_ptr;
            float sum = 0.f;
            for (int i = 0; i < K; i++) sum = __fadd_rn(sum, prob[i]);
            if (p.top_p < 1.f) {
                float c = 0.f;
                for (int i = 0; i < K; i++) {
                    c = __fadd_rn(c, __fdiv_rn(prob[i], sum));
                    if (c >= p.top_p) { n = i + 1; break; }
                }
            }
            n_p = n;
            if (p.min_p > 0.f) {
                const float cut = __fmul_rn(p.min_p, prob[0]);
                int n2 = 1;
                while (n2 < n && prob[n2] >= cut) n2++;
                n = n2;
            }
            n_m = n;
            float sum2 = 0.f;
            for (int i = 0; i < n; i++) sum2 = __fadd_rn(sum2, wgt[i]);
            const float u = __fmul_rn(rng_uniform(p.seed, (uint64_t)ctr), sum2);
            float c = 0.f;
            int pick = n - 1;
            for (int i = 0; i < n; i++) {
                c = __fadd_rn(c, wgt[i]);
                if (u < c) { pick = i; break; }
            }
            tok = key_index(sorted[pick]);
        }
        if (p.counts_out) { p.counts_out[0] = K; p.counts_out[1] = n_p; p.counts_out[2] = n_m; }
        const int step = *p.step_ptr;
        if (step < p.out_cap) p.out_tokens[step] = tok;
        *p.step_ptr = step + 1;
        *p.cur_token = tok;
        if (p.pos_ptr) *p.pos_ptr += 1;  // the sampled token is fed at the next position
    }
    if (p.cand_out && tid < K) p.cand_out[tid] = key_index(sorted[tid]);
}

__global__ void __launch_bounds__(SAMP_THREADS) sample_kernel(const SampleParams p) { sample_body(p); }

// one CTA per lane of the continuous batcher: the lane's own parameters, logits row, position and token slot
__global__ void __launch_bounds__(SAMP_THREADS) sample_batch_kernel(const SampleBatchParams b) {
    pdl_wait();
    pdl_launch_dependents();
    const int l = blockIdx.x;
    const SampleLane L = b.lanes[l];
    __shared__ int32_t step_dummy;
    if (threadIdx.x == 0) step_dummy = 0;
    __syncthreads();
    SampleParams p{};
    p.logits = b.logits + (size_t)l * b.n_vocab;
    p.n_vocab = b.n_vocab;
    p.temperature = L.temperature; p.top_k = L.top_k; p.top_p = L.top_p; p.min_p = L.min_p; p.seed = L.seed;
    p.pos_ptr = b.pos + l;
    p.step_ptr = &step_dummy;
    p.cur_token = b.tok + l;
    p.out_tokens = b.tok + l;  // step 0 of a 1-slot array: the same word
    p.out_cap = 1;
    p.mask = L.mask;
    sample_body(p);
}

void launch_sample(const SampleParams &p, cudaStream_t st) { sample_kernel<<<1, SAMP_THREADS, 0, st>>>(p); }
void launch_sample_batch(const SampleBatchParams &p, int n_lanes, cudaStream_t st) {
    launch_k(sample_batch_kernel, dim3((unsigned)n_lanes), dim3(SAMP_THREADS), 0, st, p);
}

}  // namespace aios
