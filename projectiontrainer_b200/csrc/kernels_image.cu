// Image preprocessing in front of the vision tower: decoded RGB uint8 HWC image ->
// Pillow's antialiased bicubic resize to S x S (8-bit fixed point, two passes) -> HF SiglipImageProcessor
// rescale + normalize -> `pixel_values` f32 CHW.
//
//   ref:Stage1/train_projection_stage1.py:99-104, ref:Stage2/dataset.py:90-95
//     image = Image.open(p).convert('RGB'); image = image.resize((S, S)); processor(images=image).pixel_values
//   Pillow src/libImaging/Resample.c: precompute_coeffs, normalize_coeffs_8bpc, ImagingResampleHorizontal_8bpc,
//     ImagingResampleVertical_8bpc (22 fractional bits, each pass rounds to uint8)
//   HF image_transforms.py: rescale (float64 product -> float32), normalize ((x - mean) / std in float32)
//
// Byte work bounded by reading the source image once from HBM; the coefficient tables are computed on the
// device in double precision with explicitly unfused operations so that every weight equals Pillow's.
#include <algorithm>
#include <mutex>

#include "pt_common.cuh"

#define ST(s) static_cast<cudaStream_t>(s)

namespace pt {

constexpr int kPrecisionBits = 32 - 8 - 2;
constexpr int kChunk = 128;                  // output pixels per block in the horizontal pass
constexpr int kMaxTaps = 1024;               // filter taps per output index (down-scale factors up to ~250)

struct ResampleAxis {
    int in_size, out_size, ksize;
    double scale, filterscale, support;
};

__host__ __device__ inline ResampleAxis make_axis(int in_size, int out_size) {
    ResampleAxis a;
    a.in_size = in_size;
    a.out_size = out_size;
    a.scale = static_cast<double>(static_cast<float>(in_size) - 0.0f) / out_size;
    a.filterscale = a.scale < 1.0 ? 1.0 : a.scale;
    a.support = 2.0 * a.filterscale;         // bicubic support 2.0
    a.ksize = static_cast<int>(ceil(a.support)) * 2 + 1;
    return a;
}

// Resample.c bicubic_filter (a = -0.5), same operation order, no FMA contraction
__device__ inline double bicubic_weight(double x) {
    x = fabs(x);
    if (x < 1.0)
        return __dadd_rn(__dmul_rn(__dmul_rn(__dadd_rn(__dmul_rn(1.5, x), -2.5), x), x), 1.0);
    if (x < 2.0)
        return __dmul_rn(__dadd_rn(__dmul_rn(__dadd_rn(__dmul_rn(__dadd_rn(x, -5.0), x), 8.0), x), -4.0), -0.5);
    return 0.0;
}

// One warp per output index, both axes in one launch (blockIdx.y): bounds[xx] = (first source index, tap
// count) and the fixed-point weights, stored TRANSPOSED (coeffT[tap * out_size + xx], `kpad` taps, zero beyond
// the count) so the passes read them coalesced.  The lanes evaluate the filter in parallel; the normalising sum
// is taken by lane 0 in tap order, as Pillow's loop does, so that its rounding is the same.
__global__ void __launch_bounds__(128)
resample_coeffs_kernel(ResampleAxis ax0, int2* __restrict__ bounds0, int* __restrict__ coeffT0, int kpad0,
                       ResampleAxis ax1, int2* __restrict__ bounds1, int* __restrict__ coeffT1, int kpad1) {
    __shared__ double wsh[4][kMaxTaps];
    const bool second = blockIdx.y != 0;
    const ResampleAxis ax = second ? ax1 : ax0;
    int2* bounds = second ? bounds1 : bounds0;
    int* coeffT = second ? coeffT1 : coeffT0;
    const int kpad = second ? kpad1 : kpad0;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int xx = blockIdx.x * 4 + warp;
    if (xx >= ax.out_size) return;
    const double center = __dadd_rn(0.0, __dmul_rn(xx + 0.5, ax.scale));
    const double ss = __ddiv_rn(1.0, ax.filterscale);
    int xmin = static_cast<int>(__dadd_rn(__dadd_rn(center, -ax.support), 0.5));
    if (xmin < 0) xmin = 0;
    int xmax = static_cast<int>(__dadd_rn(__dadd_rn(center, ax.support), 0.5));
    if (xmax > ax.in_size) xmax = ax.in_size;
    xmax -= xmin;
    double* w = wsh[warp];
    for (int x = lane; x < xmax; x += 32) {
        const double arg = __dmul_rn(__dadd_rn(__dadd_rn(static_cast<double>(x + xmin), -center), 0.5), ss);
        w[x] = bicubic_weight(arg);
    }
    __syncwarp();
    double ww = 0.0;
    if (lane == 0)
        for (int x = 0; x < xmax; ++x) ww = __dadd_rn(ww, w[x]);
    ww = __shfl_sync(0xffffffffu, ww, 0);
    for (int x = lane; x < kpad; x += 32) {
        int fixed = 0;
        if (x < xmax) {
            double v = w[x];
            if (ww != 0.0) v = __ddiv_rn(v, ww);
            const double scaled = __dmul_rn(v, static_cast<double>(1 << kPrecisionBits));
            fixed = static_cast<int>(v < 0 ? __dadd_rn(-0.5, scaled) : __dadd_rn(0.5, scaled));
        }
        coeffT[static_cast<long long>(x) * ax.out_size + xx] = fixed;
    }
    if (lane == 0) bounds[xx] = make_int2(xmin, xmax);
}

__device__ inline int clip8(int acc) {
    const int v = acc >> kPrecisionBits;
    return v < 0 ? 0 : (v > 255 ? 255 : v);
}

// Horizontal pass: block = (chunk of kChunk output columns, one source row).  The source bytes the chunk needs
// are staged in shared memory with aligned 4-byte loads; each thread then produces one RGB output pixel.  The
// taps of a pixel are contiguous bytes, so they are consumed four at a time: 12 bytes = three 32-bit
// shared-memory words (plus the word carried over), realigned with funnel shifts -- a quarter of the
// shared-memory wavefronts of byte loads, which is what bounded the first version.  Weights beyond a pixel's
// tap count are zero (kpad is a multiple of 4), so the ragged tail needs no branch.
__global__ void __launch_bounds__(kChunk)
resample_horizontal_kernel(const uint8_t* __restrict__ src, long long src_stride, int row0,
                           const int2* __restrict__ bounds, const int* __restrict__ coeffT, int out_w, int kpad,
                           uint8_t* __restrict__ tmp) {
    extern __shared__ __align__(16) uint8_t smem[];
    const int xx0 = blockIdx.x * kChunk;
    const int xx_last = min(xx0 + kChunk, out_w) - 1;
    const int row = blockIdx.y;
    const int2 b_first = bounds[xx0], b_last = bounds[xx_last];
    const int x_lo = b_first.x;
    const int nbytes = (b_last.x + b_last.y - x_lo) * 3;
    const uint8_t* g = src + static_cast<long long>(row0 + row) * src_stride + static_cast<long long>(x_lo) * 3;
    const int mis = static_cast<int>(reinterpret_cast<uintptr_t>(g) & 3);
    uint8_t* s = smem + mis;                               // s and g share their alignment modulo 4
    const int head = mis ? min(4 - mis, nbytes) : 0;
    const int words = (nbytes - head) >> 2;
    const int tail0 = head + (words << 2);
    if (threadIdx.x < head) s[threadIdx.x] = g[threadIdx.x];
    for (int w = threadIdx.x; w < words; w += kChunk)
        *reinterpret_cast<uint32_t*>(s + head + (w << 2)) = *reinterpret_cast<const uint32_t*>(g + head + (w << 2));
    if (threadIdx.x < nbytes - tail0) s[tail0 + threadIdx.x] = g[tail0 + threadIdx.x];
    __syncthreads();
    const int xx = xx0 + threadIdx.x;
    if (xx >= out_w) return;
    const int2 b = bounds[xx];
    const int off = mis + (b.x - x_lo) * 3;                // byte offset of the pixel's first tap inside smem
    const uint32_t* wp = reinterpret_cast<const uint32_t*>(smem) + (off >> 2);
    const int sh = (off & 3) * 8;
    int r = 1 << (kPrecisionBits - 1), gg = r, bb = r;
    uint32_t w0 = wp[0];
    const int* kp = coeffT + xx;
    for (int x = 0; x < b.y; x += 4) {
        const uint32_t w1 = wp[1], w2 = wp[2], w3 = wp[3];
        const uint32_t x0 = __funnelshift_r(w0, w1, sh), x1 = __funnelshift_r(w1, w2, sh),
                       x2 = __funnelshift_r(w2, w3, sh);          // 12 tap bytes, aligned
        const int k0 = kp[0], k1 = kp[out_w], k2 = kp[2 * static_cast<long long>(out_w)],
                  k3 = kp[3 * static_cast<long long>(out_w)];
        r += static_cast<int>(x0 & 0xff) * k0 + static_cast<int>(x0 >> 24) * k1 +
             static_cast<int>((x1 >> 16) & 0xff) * k2 + static_cast<int>((x2 >> 8) & 0xff) * k3;
        gg += static_cast<int>((x0 >> 8) & 0xff) * k0 + static_cast<int>(x1 & 0xff) * k1 +
              static_cast<int>(x1 >> 24) * k2 + static_cast<int>((x2 >> 16) & 0xff) * k3;
        bb += static_cast<int>((x0 >> 16) & 0xff) * k0 + static_cast<int>((x1 >> 8) & 0xff) * k1 +
              static_cast<int>(x2 & 0xff) * k2 + static_cast<int>(x2 >> 24) * k3;
        w0 = w3;
        wp += 3;
        kp += 4 * static_cast<long long>(out_w);
    }
    uint8_t* o = tmp + (static_cast<long long>(row) * out_w + xx) * 3;
    o[0] = static_cast<uint8_t>(clip8(r));
    o[1] = static_cast<uint8_t>(clip8(gg));
    o[2] = static_cast<uint8_t>(clip8(bb));
}

// Vertical pass + HF rescale / normalize: thread = one output pixel (3 channels), block = 128 columns of one
// output row; tmp rows are read coalesced (96 contiguous bytes per warp), the three planes are written coalesced.
__global__ void __launch_bounds__(128)
resample_vertical_normalize_kernel(const uint8_t* __restrict__ tmp, int row_shift, const int2* __restrict__ bounds,
                                   const int* __restrict__ coeffT, int out_h, int out_w, double rescale,
                                   float m0, float m1, float m2, float s0, float s1, float s2,
                                   float* __restrict__ dst, uint8_t* __restrict__ dst_u8) {
    const int xx = blockIdx.x * 128 + threadIdx.x;
    const int yy = blockIdx.y;
    if (xx >= out_w) return;
    const int2 b = bounds[yy];
    const uint8_t* p = tmp + (static_cast<long long>(b.x - row_shift) * out_w + xx) * 3;
    const long long rs = static_cast<long long>(out_w) * 3;
    int r = 1 << (kPrecisionBits - 1), g = r, bl = r;
    for (int y = 0; y < b.y; ++y) {
        const int k = coeffT[static_cast<long long>(y) * out_h + yy];
        r += p[0] * k;
        g += p[1] * k;
        bl += p[2] * k;
        p += rs;
    }
    const int v[3] = {clip8(r), clip8(g), clip8(bl)};
    const float mean[3] = {m0, m1, m2}, sd[3] = {s0, s1, s2};
    const long long plane = static_cast<long long>(out_h) * out_w;
    const long long o = static_cast<long long>(yy) * out_w + xx;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        // rescale: float64 product cast to float32; normalize: (x - mean) / std in float32, both unfused
        const float x = __double2float_rn(__dmul_rn(static_cast<double>(v[c]), rescale));
        dst[c * plane + o] = __fdiv_rn(__fsub_rn(x, mean[c]), sd[c]);
        if (dst_u8) dst_u8[o * 3 + c] = static_cast<uint8_t>(v[c]);
    }
}

struct ImagePlan {
    ResampleAxis h, v;            // horizontal (over W), vertical (over H)
    int kpad_h, kpad_v;           // taps stored per output index (multiple of 4, zero beyond the count)
    long long off_bounds_h, off_bounds_v, off_coeff_h, off_coeff_v, off_tmp, total;
    int smem_bytes;
};

static ImagePlan plan_image(int H, int W, int out_h, int out_w) {
    ImagePlan p;
    p.h = make_axis(W, out_w);
    p.v = make_axis(H, out_h);
    auto up = [](long long x) { return (x + 255) / 256 * 256; };
    p.kpad_h = (p.h.ksize + 3) / 4 * 4;
    p.kpad_v = (p.v.ksize + 3) / 4 * 4;
    long long o = 0;
    p.off_bounds_h = o; o = up(o + static_cast<long long>(out_w) * sizeof(int2));
    p.off_bounds_v = o; o = up(o + static_cast<long long>(out_h) * sizeof(int2));
    p.off_coeff_h = o;  o = up(o + static_cast<long long>(out_w) * p.kpad_h * sizeof(int));
    p.off_coeff_v = o;  o = up(o + static_cast<long long>(out_h) * p.kpad_v * sizeof(int));
    p.off_tmp = o;      o = up(o + static_cast<long long>(H) * out_w * 3);
    p.total = o;
    // widest source span of one chunk: kChunk centres apart by `scale`, plus the filter support on both sides
    const long long span = static_cast<long long>(ceil(kChunk * p.h.scale + 2 * p.h.support)) + 4;
    p.smem_bytes = static_cast<int>(std::min<long long>(span, W) * 3 + 8 + 32);      // + over-read of the last 4-tap group
    return p;
}

}  // namespace pt

using namespace pt;

extern "C" long long pt_image_preprocess_workspace(int H, int W, int out_h, int out_w) {
    if (H <= 0 || W <= 0 || out_h <= 0 || out_w <= 0) return -1;
    return plan_image(H, W, out_h, out_w).total;
}

extern "C" int pt_image_preprocess(const uint8_t* src_rgb, int H, int W, long long src_row_stride, float* dst_chw,
                                   uint8_t* dst_u8_hwc, int out_h, int out_w, float mean0, float mean1, float mean2,
                                   float std0, float std1, float std2, double rescale, void* workspace,
                                   long long workspace_bytes, void* stream) {
    PT_REQUIRE(src_rgb && dst_chw && workspace, "pt_image_preprocess: null pointer");
    PT_REQUIRE(H > 0 && W > 0 && out_h > 0 && out_w > 0 && H <= 65535 && W < (1 << 24) && out_h <= 65535,
               "pt_image_preprocess: bad image size (rows are a grid dimension: at most 65535)");
    PT_REQUIRE(src_row_stride >= static_cast<long long>(W) * 3, "pt_image_preprocess: row stride shorter than a row");
    PT_REQUIRE(std0 != 0.f && std1 != 0.f && std2 != 0.f, "pt_image_preprocess: zero std");
    PT_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 15) == 0, "pt_image_preprocess: workspace must be 16B aligned");
    const ImagePlan p = plan_image(H, W, out_h, out_w);
    PT_REQUIRE(workspace_bytes >= p.total, "pt_image_preprocess: workspace too small (see pt_image_preprocess_workspace)");
    PT_REQUIRE(p.smem_bytes <= 200 * 1024, "pt_image_preprocess: down-scale factor too large for the shared-memory stage");
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    cudaStream_t st = ST(stream);
    uint8_t* ws = static_cast<uint8_t*>(workspace);
    int2* bounds_h = reinterpret_cast<int2*>(ws + p.off_bounds_h);
    int2* bounds_v = reinterpret_cast<int2*>(ws + p.off_bounds_v);
    int* coeff_h = reinterpret_cast<int*>(ws + p.off_coeff_h);
    int* coeff_v = reinterpret_cast<int*>(ws + p.off_coeff_v);
    uint8_t* tmp = ws + p.off_tmp;
    PT_REQUIRE(p.h.ksize <= kMaxTaps && p.v.ksize <= kMaxTaps, "pt_image_preprocess: down-scale factor too large");
    dim3 gc((std::max(out_w, out_h) + 3) / 4, 2);
    resample_coeffs_kernel<<<gc, 128, 0, st>>>(p.h, bounds_h, coeff_h, p.kpad_h, p.v, bounds_v, coeff_v, p.kpad_v);
    // The horizontal pass covers every source row: which rows the vertical pass reads is only known on the
    // device (its bounds), and the rows outside that range are at most the filter support at either end.
    if (p.smem_bytes > 48 * 1024) {
        if (cudaFuncSetAttribute(resample_horizontal_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 p.smem_bytes) != cudaSuccess)
            return check_launch("image horizontal smem attribute");
    }
    dim3 gh((out_w + kChunk - 1) / kChunk, H);
    resample_horizontal_kernel<<<gh, kChunk, p.smem_bytes, st>>>(src_rgb, src_row_stride, 0, bounds_h, coeff_h, out_w,
                                                                p.kpad_h, tmp);
    dim3 gv((out_w + 127) / 128, out_h);
    resample_vertical_normalize_kernel<<<gv, 128, 0, st>>>(tmp, 0, bounds_v, coeff_v, out_h, out_w, rescale, mean0,
                                                           mean1, mean2, std0, std1, std2, dst_chw, dst_u8_hwc);
    return check_launch("image_preprocess");
}

// ---- allocation-free variant ------------------------------------------------------------------------------------
// pt_image_preprocess_static: the same three kernels with their scratch (tap tables + the horizontally resampled
// intermediate) in a file-scope device array instead of a caller-allocated workspace, for callers that cannot
// allocate: code under CUDA graph capture, serving loops, harnesses that forbid cudaMalloc.  No counterpart in the
// reference.  Every GPU that loads the library holds the array until
// exit.  All calls on a device share it: calls that overlap on the GPU (different streams) race.  Nothing in it has
// to start at zero — every byte a call reads (bounds, taps including their zero padding, intermediate rows) was
// written earlier in the same call — so the previous call's contents need no clearing.
namespace pt {
constexpr long long kImageStaticScratchBytes = 8ll << 20;      // 8 MiB: e.g. a 6900-row source down to 384 x 384
static __device__ __align__(256) uint8_t g_image_scratch[kImageStaticScratchBytes];

// Device address of g_image_scratch on the CURRENT device (the host-side symbol is only a handle), looked up once
// per device.
static int image_static_scratch(void** out) {
    constexpr int kMaxDev = 64;
    static void* ptrs[kMaxDev] = {};
    static std::mutex mu;
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDev) {
        set_error("pt_image_preprocess_static: cudaGetDevice failed or device ordinal >= %d", kMaxDev);
        return PT_ERR_CUDA;
    }
    std::lock_guard<std::mutex> lock(mu);
    if (ptrs[dev] == nullptr) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_image_scratch);
        if (e != cudaSuccess || p == nullptr) {
            set_error("pt_image_preprocess_static: cudaGetSymbolAddress failed on device %d: %s", dev,
                      cudaGetErrorString(e));
            return PT_ERR_CUDA;
        }
        ptrs[dev] = p;
    }
    *out = ptrs[dev];
    return PT_OK;
}
}  // namespace pt

extern "C" long long pt_image_preprocess_static_max_bytes(void) { return pt::kImageStaticScratchBytes; }

extern "C" int pt_image_preprocess_static(const uint8_t* src_rgb, int H, int W, long long src_row_stride,
                                          float* dst_chw, uint8_t* dst_u8_hwc, int out_h, int out_w, float mean0,
                                          float mean1, float mean2, float std0, float std1, float std2,
                                          double rescale, void* stream) {
    PT_REQUIRE(H > 0 && W > 0 && out_h > 0 && out_w > 0, "pt_image_preprocess_static: bad image size");
    const long long need = plan_image(H, W, out_h, out_w).total;
    PT_REQUIRE(need <= kImageStaticScratchBytes,
               "pt_image_preprocess_static: a %d x %d -> %d x %d resize needs %lld bytes of scratch, the built-in "
               "array holds %lld: use pt_image_preprocess with a caller-allocated workspace", H, W, out_h, out_w,
               need, kImageStaticScratchBytes);
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    void* ws = nullptr;
    const int rc = image_static_scratch(&ws);
    if (rc != PT_OK) return rc;
    return pt_image_preprocess(src_rgb, H, W, src_row_stride, dst_chw, dst_u8_hwc, out_h, out_w, mean0, mean1, mean2,
                               std0, std1, std2, rescale, ws, kImageStaticScratchBytes, stream);
}
