// Next-token choice of `generate` on the device, one launch per generated token (replaces, per token, ~25 torch
// launches over the [B, V] logits: clone, gather / scatter of the repetition penalty, top-k, a full sort of V = 262 144
// values for top-p, two softmaxes, cumsum, scatter, multinomial, where, the bookkeeping of finished rows — 2 ms per token
// at B = 64, more than the decode step itself).
//
// HF's order (HF:generation/utils.py `_get_logits_processor` + `_sample`, called at ref:Stage1/inference_stage1.py:176-186
// with do_sample, temperature 0.7, top_p 0.9 and HF's default top_k = 50):
//   repetition penalty (x < 0 ? x * rp : x / rp on the tokens generated so far) -> / temperature -> top-k (everything
//   below the k-th largest value becomes -inf; ties with it stay) -> top-p (ascending order, tokens whose cumulative
//   probability is <= 1 - top_p go; the largest stays) -> softmax -> one multinomial draw;  or argmax without sampling;
//   then: finished rows emit pad, out[:, step] = token, finished |= token == eos.
// The draw is the inverse CDF over the surviving tokens in descending order at u[b] (a uniform number the caller
// generates): the same distribution as torch.multinomial, not the same random stream.
//
// One CTA of 1024 threads per row.  Pass 1 reads the row once (strided, so that a thread's 256 values are a sample of
// the row) and keeps each thread's maximum; the k-th largest of the 1024 maxima is a lower bound of the k-th largest
// value, and pass 2 (the row is in L2 by now) collects everything at or above it: ~k .. 2k candidates for any row that
// is not mostly ties.  The candidates are ranked by counting, which gives the exact k-th value and the descending
// order at once.  If more than kCap values pass the bound (flat rows), the bound is raised by bisection on the ordered
// integer image of the floats: at most 32 more passes, exact.
#include "pt_common.cuh"

namespace pt {

constexpr int kSampThreads = 1024;
constexpr int kCap = 2048;                              // candidates held in shared memory

__device__ __forceinline__ uint32_t fkey(float x) {     // order-preserving float -> uint32
    const uint32_t u = __float_as_uint(x);
    return u ^ ((u >> 31) ? 0xffffffffu : 0x80000000u);
}

struct SampleArgs {
    float* logits;                  // [B, ld]; the repetition penalty is applied to it in place
    long long ld;
    int V;
    const long long* prev;          // [B, ld_prev]: tokens generated so far (repetition penalty), or null
    long long ld_prev;
    int nprev;
    float rep, inv_temp, top_p;
    int top_k;                      // 0: argmax (no sampling)
    const float* u;                 // [B] uniform in [0, 1)
    uint8_t* done;                  // [B] in/out, or null
    long long pad_id, eos_id;       // < 0: none
    long long* out;                 // out[b * ld_out] = token
    long long ld_out;
    long long* next;                // [B] the token again (the decode step's input), or null
};

__global__ void __launch_bounds__(kSampThreads) sample_token_kernel(const SampleArgs p) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    __shared__ float s_max[kSampThreads];
    __shared__ int s_arg[kSampThreads];
    __shared__ float cv[kCap];                          // candidates: value, index; then sorted copies
    __shared__ int ci[kCap];
    __shared__ float sv[kCap];
    __shared__ int si[kCap];
    __shared__ int s_cnt, s_tok;
    __shared__ float s_bound;
    __shared__ uint32_t s_lo, s_hi;
    uint32_t* bitmap = reinterpret_cast<uint32_t*>(smem_raw);   // V bits: token was generated before
    const int b = blockIdx.x, tid = threadIdx.x;
    float* row = p.logits + static_cast<long long>(b) * p.ld;
    const bool pen = p.prev != nullptr && p.nprev > 0 && p.rep != 1.f;
    if (pen) {      // the penalty is applied IN PLACE, once per distinct token (the bitmap only de-duplicates)
        for (int i = tid; i < (p.V + 31) / 32; i += kSampThreads) bitmap[i] = 0u;
        __syncthreads();
        const long long* pr = p.prev + static_cast<long long>(b) * p.ld_prev;
        for (int i = tid; i < p.nprev; i += kSampThreads) {
            const long long t = pr[i];
            if (t >= 0 && t < p.V) {
                const uint32_t bit = 1u << (t & 31);
                if (!(atomicOr(&bitmap[t >> 5], bit) & bit)) {
                    const float x = row[t];
                    row[t] = x < 0.f ? x * p.rep : x / p.rep;
                }
            }
        }
        __syncthreads();
    }
    auto value = [&](int i) {
        const float x = row[i];
        return x != x ? -INFINITY : x;                  // a NaN logit never wins
    };
    // every (index, value) of the row through fn: 16-byte loads, four per thread in flight
    auto fix = [&](int, float x) { return x != x ? -INFINITY : x; };
    const int n4 = reinterpret_cast<uintptr_t>(row) % 16 == 0 ? p.V / 4 : 0;
    auto scan = [&](auto&& fn) {
        constexpr int U = 4;
        const float4* row4 = reinterpret_cast<const float4*>(row);
        for (int j = tid; j < n4; j += kSampThreads * U) {
            float4 v[U];
#pragma unroll
            for (int u = 0; u < U; ++u)
                if (j + u * kSampThreads < n4) v[u] = row4[j + u * kSampThreads];   // (plain loads: the row was just patched)
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int i = 4 * (j + u * kSampThreads);
                if (i < 4 * n4) {
                    fn(i, fix(i, v[u].x)); fn(i + 1, fix(i + 1, v[u].y));
                    fn(i + 2, fix(i + 2, v[u].z)); fn(i + 3, fix(i + 3, v[u].w));
                }
            }
        }
        for (int i = 4 * n4 + tid; i < p.V; i += kSampThreads) fn(i, fix(i, row[i]));
    };

    // pass 1: per-thread maximum (smallest index on ties)
    float m = -INFINITY;
    int am = p.V;
    scan([&](int i, float x) {
        if (x > m || (x == m && i < am)) { m = x; am = i; }
    });
    s_max[tid] = m;
    s_arg[tid] = am;
    __syncthreads();
    int tok;
    if (p.top_k <= 0) {                                 // argmax: the largest value, the smallest index among ties
        for (int s = kSampThreads / 2; s > 0; s >>= 1) {
            if (tid < s) {
                const float o = s_max[tid + s];
                const int oa = s_arg[tid + s];
                if (o > s_max[tid] || (o == s_max[tid] && oa < s_arg[tid])) { s_max[tid] = o; s_arg[tid] = oa; }
            }
            __syncthreads();
        }
        tok = s_arg[0] < p.V ? s_arg[0] : 0;
    } else {
        const int k = p.top_k < p.V ? p.top_k : p.V;
        // the k-th largest thread maximum is <= the k-th largest value (k <= 1024; beyond that: no bound)
        if (tid == 0) { s_bound = -INFINITY; s_cnt = 0; }
        __syncthreads();
        if (k <= kSampThreads) {    // radix select over the 1024 ordered keys, 8 bits a pass
            uint32_t* hist = reinterpret_cast<uint32_t*>(sv);    // 256 bins (sv is free until the ranking below)
            const uint32_t key = fkey(m);
            uint32_t prefix = 0;
            int want = k;
#pragma unroll 1
            for (int pass = 0; pass < 4; ++pass) {
                const int shift = 24 - 8 * pass;
                if (tid < 256) hist[tid] = 0u;
                __syncthreads();
                if (pass == 0 || (key >> (shift + 8)) == prefix) atomicAdd(&hist[(key >> shift) & 255u], 1u);
                __syncthreads();
                if (tid == 0) {                          // the bin that holds the want-th largest of the matching keys
                    int above = 0, bin = 255;
                    for (; bin > 0; --bin) {
                        const int c = static_cast<int>(hist[bin]);
                        if (above + c >= want) break;
                        above += c;
                    }
                    s_lo = static_cast<uint32_t>(bin);
                    s_hi = static_cast<uint32_t>(above);
                }
                __syncthreads();
                prefix = (prefix << 8) | s_lo;
                want -= static_cast<int>(s_hi);
                __syncthreads();
            }
            if (tid == 0) {
                const uint32_t u = prefix ^ ((prefix >> 31) ? 0x80000000u : 0xffffffffu);
                s_bound = __uint_as_float(u);
            }
        }
        __syncthreads();
        // pass 2: collect everything >= the bound; raise the bound by bisection while it lets too much through
        auto collect = [&](float bound) {
            scan([&](int i, float x) {
                if (x >= bound) {
                    const int at = atomicAdd(&s_cnt, 1);
                    if (at < kCap) { cv[at] = x; ci[at] = i; }
                }
            });
            __syncthreads();
        };
        collect(s_bound);
        if (s_cnt > kCap) {
            if (tid == 0) { s_lo = fkey(s_bound); s_hi = 0xffffffffu; }   // count(>= lo) > cap; count(>= hi) taken as < k
            __syncthreads();
            for (int it = 0; it < 34 && s_cnt > kCap; ++it) {
                const uint32_t lo = s_lo, hi = s_hi;
                if (hi - lo <= 1u) break;
                const uint32_t mid = lo + (hi - lo) / 2;
                __syncthreads();
                if (tid == 0) s_cnt = 0;
                __syncthreads();
                int c = 0;
                scan([&](int, float x) { c += fkey(x) >= mid; });
                atomicAdd(&s_cnt, c);
                __syncthreads();
                const int tot = s_cnt;
                __syncthreads();
                if (tid == 0) {
                    if (tot >= k) s_lo = mid; else s_hi = mid;      // the k-th largest key is >= mid / < mid
                    s_cnt = tot >= k && tot <= kCap ? 0 : kCap + 1;  // done when k <= count <= cap
                }
                __syncthreads();
                if (s_cnt == 0) {                         // collect at this bound
                    const uint32_t bk = s_lo;
                    scan([&](int i, float x) {
                        if (fkey(x) >= bk) {
                            const int at = atomicAdd(&s_cnt, 1);
                            if (at < kCap) { cv[at] = x; ci[at] = i; }
                        }
                    });
                    __syncthreads();
                    break;
                }
            }
            if (s_cnt > kCap || s_cnt == 0) {            // more than kCap ties at the k-th value: the lowest indices
                __syncthreads();                         // of the tied block that fit (a flat row; any choice among
                if (tid == 0) s_cnt = 0;                 // equal logits has the same probability)
                __syncthreads();
                const uint32_t bk = s_lo;
                scan([&](int i, float x) {
                    if (fkey(x) > bk) {
                        const int at = atomicAdd(&s_cnt, 1);
                        if (at < kCap) { cv[at] = x; ci[at] = i; }
                    }
                });
                __syncthreads();
                if (tid == 0) {
                    int n = s_cnt < kCap ? s_cnt : kCap;
                    for (int i = 0; i < p.V && n < kCap; ++i)
                        if (fkey(value(i)) == bk) { cv[n] = value(i); ci[n] = i; ++n; }
                    s_cnt = n;
                }
                __syncthreads();
            }
        }
        const int n = s_cnt < kCap ? s_cnt : kCap;
        // rank the candidates (descending value, ascending index): exact k-th value and the sorted order
        for (int i = tid; i < n; i += kSampThreads) {
            const float x = cv[i];
            const int xi = ci[i];
            int rank = 0;
            for (int j = 0; j < n; ++j) rank += (cv[j] > x) || (cv[j] == x && ci[j] < xi);
            sv[rank] = x;
            si[rank] = xi;
        }
        __syncthreads();
        if (tid == 0) {
            const int kk = k < n ? k : n;
            const float kth = sv[kk - 1];
            int keep = kk;
            while (keep < n && sv[keep] == kth) ++keep;   // ties with the k-th value stay (HF masks `< kth`)
            // softmax over the kept values at the temperature, in descending order
            const float mx = sv[0];
            float Z = 0.f;
            for (int i = 0; i < keep; ++i) { cv[i] = __expf((sv[i] - mx) * p.inv_temp); Z += cv[i]; }
            // top-p: a token stays while the probability above it is < top_p (the first always)
            int nk = keep;
            if (p.top_p < 1.f) {
                float above = 0.f;
                nk = 0;
                for (int i = 0; i < keep; ++i) {
                    if (i > 0 && !(above < p.top_p * Z)) break;
                    above += cv[i];
                    ++nk;
                }
            }
            float Zk = 0.f;
            for (int i = 0; i < nk; ++i) Zk += cv[i];
            const float target = p.u[b] * Zk;
            float acc = 0.f;
            int pick = nk - 1;
            for (int i = 0; i < nk; ++i) {
                acc += cv[i];
                if (target < acc) { pick = i; break; }
            }
            s_tok = si[pick];
        }
        __syncthreads();
        tok = s_tok;
    }
    if (tid == 0) {
        long long t = tok;
        const bool was_done = p.done != nullptr && p.done[b] != 0;
        if (was_done && p.pad_id >= 0) t = p.pad_id;
        p.out[static_cast<long long>(b) * p.ld_out] = t;
        if (p.next) p.next[b] = t;
        if (p.done != nullptr && p.eos_id >= 0 && t == p.eos_id) p.done[b] = 1;
    }
}

}  // namespace pt

using namespace pt;

extern "C" int pt_sample_token(float* logits, long long ld, int B, int V, const long long* prev, long long ld_prev,
                               int nprev, float repetition_penalty, float temperature, int top_k, float top_p,
                               const float* u, uint8_t* done, long long pad_id, long long eos_id, long long* out,
                               long long ld_out, long long* next, void* stream) {
    PT_REQUIRE(logits && out && B > 0 && V > 0 && ld >= V, "pt_sample_token: bad arguments");
    PT_REQUIRE(top_k >= 0 && top_k <= kSampThreads, "pt_sample_token: top_k must be 0 (argmax) .. 1024");
    PT_REQUIRE(top_k == 0 || (u != nullptr && temperature > 0.f && top_p > 0.f),
               "pt_sample_token: sampling needs u, temperature > 0 and top_p > 0");
    PT_REQUIRE(repetition_penalty > 0.f, "pt_sample_token: repetition_penalty must be > 0");
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    const bool pen = prev != nullptr && nprev > 0 && repetition_penalty != 1.f;
    const size_t smem = pen ? static_cast<size_t>((V + 31) / 32) * 4 : 0;
    PT_REQUIRE(smem <= 160 * 1024, "pt_sample_token: vocabulary too large for the repetition-penalty bitmap");
    if (smem > 0) {
        static bool done_attr[64] = {};
        int dev = 0;
        cudaGetDevice(&dev);
        if (dev < 0 || dev >= 64 || !done_attr[dev]) {
            const cudaError_t e = cudaFuncSetAttribute(sample_token_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                       160 * 1024);
            if (e != cudaSuccess) { set_error("cudaFuncSetAttribute(sample_token): %s", cudaGetErrorString(e)); return PT_ERR_CUDA; }
            if (dev >= 0 && dev < 64) done_attr[dev] = true;
        }
    }
    SampleArgs a;
    a.logits = logits; a.ld = ld; a.V = V; a.prev = prev; a.ld_prev = ld_prev; a.nprev = nprev;
    a.rep = repetition_penalty; a.inv_temp = top_k > 0 ? 1.f / temperature : 1.f; a.top_p = top_p; a.top_k = top_k;
    a.u = u; a.done = done; a.pad_id = pad_id; a.eos_id = eos_id; a.out = out; a.ld_out = ld_out; a.next = next;
    sample_token_kernel<<<B, kSampThreads, smem, static_cast<cudaStream_t>(stream)>>>(a);
    return check_launch("sample_token");
}
