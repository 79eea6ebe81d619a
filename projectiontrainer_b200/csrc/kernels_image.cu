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

#include "pt_common.cuh"

#define ST(s) static_cast<cudaStream_t>(s)

namespace pt {

constexpr int kPrecisionBits = 32 - 8 - 2;
constexpr int kChunk = 128;                  // output pixels per block in the horizontal pass

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

// One thread per output index of one axis: bounds[xx] = (first source index, tap count), and the fixed-point
// weights, stored TRANSPOSED (coeffT[tap * out_size + xx]) so the passes read them coalesced.
__global__ void resample_coeffs_kernel(ResampleAxis ax, int2* __restrict__ bounds, int* __restrict__ coeffT) {
    const int xx = blockIdx.x * blockDim.x + threadIdx.x;
    if (xx >= ax.out_size) return;
    const double center = __dadd_rn(0.0, __dmul_rn(xx + 0.5, ax.scale));
    const double ss = __ddiv_rn(1.0, ax.filterscale);
    int xmin = static_cast<int>(__dadd_rn(__dadd_rn(center, -ax.support), 0.5));
    if (xmin < 0) xmin = 0;
    int xmax = static_cast<int>(__dadd_rn(__dadd_rn(center, ax.support), 0.5));
    if (xmax > ax.in_size) xmax = ax.in_size;
    xmax -= xmin;
    double ww = 0.0;
    for (int x = 0; x < xmax; ++x) {
        const double arg = __dmul_rn(__dadd_rn(__dadd_rn(static_cast<double>(x + xmin), -center), 0.5), ss);
        ww = __dadd_rn(ww, bicubic_weight(arg));
    }
    for (int x = 0; x < ax.ksize; ++x) {
        int fixed = 0;
        if (x < xmax) {
            const double arg = __dmul_rn(__dadd_rn(__dadd_rn(static_cast<double>(x + xmin), -center), 0.5), ss);
            double w = bicubic_weight(arg);
            if (ww != 0.0) w = __ddiv_rn(w, ww);
            const double scaled = __dmul_rn(w, static_cast<double>(1 << kPrecisionBits));
            fixed = static_cast<int>(w < 0 ? __dadd_rn(-0.5, scaled) : __dadd_rn(0.5, scaled));
        }
        coeffT[static_cast<long long>(x) * ax.out_size + xx] = fixed;
    }
    bounds[xx] = make_int2(xmin, xmax);
}

__device__ inline int clip8(int acc) {
    const int v = acc >> kPrecisionBits;
    return v < 0 ? 0 : (v > 255 ? 255 : v);
}

// Horizontal pass: block = (chunk of kChunk output columns, one source row).  The source bytes the chunk needs
// are staged in shared memory with aligned 4-byte loads; each thread then produces one RGB output pixel.
__global__ void __launch_bounds__(kChunk)
resample_horizontal_kernel(const uint8_t* __restrict__ src, long long src_stride, int row0,
                           const int2* __restrict__ bounds, const int* __restrict__ coeffT, int out_w, int ksize,
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
    const uint8_t* p = s + (b.x - x_lo) * 3;
    int r = 1 << (kPrecisionBits - 1), gg = r, bb = r;
    for (int x = 0; x < b.y; ++x) {
        const int k = coeffT[static_cast<long long>(x) * out_w + xx];
        r += p[0] * k;
        gg += p[1] * k;
        bb += p[2] * k;
        p += 3;
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
    long long off_bounds_h, off_bounds_v, off_coeff_h, off_coeff_v, off_tmp, total;
    int smem_bytes;
};

static ImagePlan plan_image(int H, int W, int out_h, int out_w) {
    ImagePlan p;
    p.h = make_axis(W, out_w);
    p.v = make_axis(H, out_h);
    auto up = [](long long x) { return (x + 255) / 256 * 256; };
    long long o = 0;
    p.off_bounds_h = o; o = up(o + static_cast<long long>(out_w) * sizeof(int2));
    p.off_bounds_v = o; o = up(o + static_cast<long long>(out_h) * sizeof(int2));
    p.off_coeff_h = o;  o = up(o + static_cast<long long>(out_w) * p.h.ksize * sizeof(int));
    p.off_coeff_v = o;  o = up(o + static_cast<long long>(out_h) * p.v.ksize * sizeof(int));
    p.off_tmp = o;      o = up(o + static_cast<long long>(H) * out_w * 3);
    p.total = o;
    // widest source span of one chunk: kChunk centres apart by `scale`, plus the filter support on both sides
    const long long span = static_cast<long long>(ceil(kChunk * p.h.scale + 2 * p.h.support)) + 4;
    p.smem_bytes = static_cast<int>(std::min<long long>(span, W) * 3 + 8);
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
    resample_coeffs_kernel<<<(out_w + 127) / 128, 128, 0, st>>>(p.h, bounds_h, coeff_h);
    resample_coeffs_kernel<<<(out_h + 127) / 128, 128, 0, st>>>(p.v, bounds_v, coeff_v);
    // The horizontal pass covers every source row: which rows the vertical pass reads is only known on the
    // device (its bounds), and the rows outside that range are at most the filter support at either end.
    if (p.smem_bytes > 48 * 1024) {
        if (cudaFuncSetAttribute(resample_horizontal_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 p.smem_bytes) != cudaSuccess)
            return check_launch("image horizontal smem attribute");
    }
    dim3 gh((out_w + kChunk - 1) / kChunk, H);
    resample_horizontal_kernel<<<gh, kChunk, p.smem_bytes, st>>>(src_rgb, src_row_stride, 0, bounds_h, coeff_h, out_w,
                                                                p.h.ksize, tmp);
    dim3 gv((out_w + 127) / 128, out_h);
    resample_vertical_normalize_kernel<<<gv, 128, 0, st>>>(tmp, 0, bounds_v, coeff_v, out_h, out_w, rescale, mean0,
                                                           mean1, mean2, std0, std1, std2, dst_chw, dst_u8_hwc);
    return check_launch("image_preprocess");
}
