// K1: image pyramid for pyramidal LK.
//
// Replaces the implicit cv::buildOpticalFlowPyramid + calcScharrDeriv work inside both
// cv2.calcOpticalFlowPyrLK calls of the reference (/root/reference/src/vo_pipeline.py:215, :501;
// behaviour in SURVEY.md App. A.1 steps 1-2).  The reference rebuilds both pyramids on every call
// (4 pyramid builds per frame); here each frame is built once and cached by key.
//
// Per level, one launch produces (a) the u8 level with a kPad REFLECT_101 border, so LK windows
// that hang over the edge need no index clamping, and (b) the int16 Scharr (Ix,Iy) plane, zero
// outside the image.  Level l+1 is the integer 5-tap [1 4 6 4 1] pyrDown, (s+128)>>8.
#include "ctx.cuh"

namespace vo {

namespace {

constexpr int TX = 32, TY = 16;

__device__ __forceinline__ int reflect101(int p, int len)
{
    if (len == 1) return 0;
    while (p < 0 || p >= len) p = p < 0 ? -p : 2 * (len - 1) - p;
    return p;
}

// value of level pixel (x,y), 0<=x<w, 0<=y<h
template <bool RAW>
__device__ __forceinline__ int level_value(const uint8_t* __restrict__ src, int sstride, int x, int y)
{
    if (RAW) return src[(size_t)y * sstride + x];
    // src points at the padded origin of the finer level; taps 2x-2..2x+2 stay inside the border
    const uint8_t* p = src + (size_t)(2 * y - 2 + kPad) * sstride + (2 * x - 2 + kPad);
    int acc = 0;
#pragma unroll
    for (int j = 0; j < 5; j++) {
        const int kj = (j == 0 || j == 4) ? 1 : (j == 2 ? 6 : 4);
        const uint8_t* r = p + (size_t)j * sstride;
        acc += kj * (r[0] + 4 * r[1] + 6 * r[2] + 4 * r[3] + r[4]);
    }
    return (acc + 128) >> 8;
}

template <bool RAW>
__global__ void __launch_bounds__(TX* TY)
build_level_kernel(const uint8_t* __restrict__ src, int sstride, LevelDesc dst)
{
    __shared__ uint8_t tile[TY + 2][TX + 2 + 2];
    pdl_prologue();
    const int X0 = blockIdx.x * TX - kPad, Y0 = blockIdx.y * TY - kPad;
    const int tid = threadIdx.y * TX + threadIdx.x;
    for (int i = tid; i < (TY + 2) * (TX + 2); i += TX * TY) {
        const int sy = i / (TX + 2), sx = i - sy * (TX + 2);
        const int x = reflect101(X0 + sx - 1, dst.w), y = reflect101(Y0 + sy - 1, dst.h);
        tile[sy][sx] = (uint8_t)level_value<RAW>(src, sstride, x, y);
    }
    __syncthreads();
    const int X = X0 + threadIdx.x, Y = Y0 + threadIdx.y;
    if (X >= dst.w + kPad || Y >= dst.h + kPad) return;
    const int tx = threadIdx.x + 1, ty = threadIdx.y + 1;
    const size_t o = (size_t)(Y + kPad) * dst.stride + (X + kPad);
    dst.img[o] = tile[ty][tx];
    if (X >= 0 && X < dst.w && Y >= 0 && Y < dst.h) {
        // Scharr, gain 32: t0 = 3*(I[y-1]+I[y+1]) + 10*I[y], t1 = I[y+1]-I[y-1]
        const int t0m = 3 * (tile[ty - 1][tx - 1] + tile[ty + 1][tx - 1]) + 10 * tile[ty][tx - 1];
        const int t0p = 3 * (tile[ty - 1][tx + 1] + tile[ty + 1][tx + 1]) + 10 * tile[ty][tx + 1];
        const int t1m = tile[ty + 1][tx - 1] - tile[ty - 1][tx - 1];
        const int t1c = tile[ty + 1][tx] - tile[ty - 1][tx];
        const int t1p = tile[ty + 1][tx + 1] - tile[ty - 1][tx + 1];
        dst.deriv[o] = make_short2((short)(t0p - t0m), (short)(3 * (t1m + t1p) + 10 * t1c));
    }
}

}  // namespace

// cuTensorMapEncodeTiled through the runtime's driver entry point (the library links against cudart only)
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_tiled_fn()
{
    static EncodeTiledFn fn = []() -> EncodeTiledFn {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) {
            cudaGetLastError();
            return nullptr;
        }
        return (EncodeTiledFn)p;
    }();
    return fn;
}

// TMA descriptors of the padded u8 planes: 2-D tensors (row stride x rows) of bytes, box 64 B x 32 rows = the tile
// the KLT kernel stages per warp and level around a track's starting position (klt.cu)
static void encode_level_maps(FrameSlot& s)
{
    s.tmap_valid = false;
    s.pyr.tmaps = nullptr;
    EncodeTiledFn fn = encode_tiled_fn();
    if (!fn) return;
    for (int l = 0; l < 4; l++) {
        if (l >= s.pyr.nlevels) { memset(&s.tmap_img[l], 0, sizeof(CUtensorMap)); continue; }
        const LevelDesc& d = s.pyr.lv[l];
        const cuuint64_t gdim[2] = { (cuuint64_t)d.stride, (cuuint64_t)(d.h + 2 * kPad) };
        const cuuint64_t gstride[1] = { (cuuint64_t)d.stride };
        const cuuint32_t box[2] = { 64, 32 };
        const cuuint32_t estr[2] = { 1, 1 };
        if (fn(&s.tmap_img[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, d.img, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
               CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
            return;
    }
    s.tmap_valid = true;
    s.pyr.tmaps = s.tmap_img;
}

int build_pyramid(Ctx* c, FrameSlot& s, int max_level_cap, cudaStream_t stream)
{
    if (!stream) stream = c->stream;
    // geometry
    int w = s.w, h = s.h, nl = 0;
    size_t img_total = 0;
    int ws[kMaxLevels], hs[kMaxLevels], st[kMaxLevels];
    size_t off[kMaxLevels];
    for (int l = 0; l <= max_level_cap && l < kMaxLevels; l++) {
        ws[l] = w; hs[l] = h;
        st[l] = (w + 2 * kPad + 63) & ~63;
        off[l] = img_total;
        img_total += (size_t)st[l] * (h + 2 * kPad);
        nl = l + 1;
        if (w <= 2 || h <= 2) break;
        w = (w + 1) / 2; h = (h + 1) / 2;
    }
    if (img_total > s.img_bytes) {
        if (s.img_pool) { cudaStreamSynchronize(c->stream); cudaStreamSynchronize(stream); cudaFree(s.img_pool); cudaFree(s.deriv_pool); }
        s.img_pool = nullptr; s.deriv_pool = nullptr; s.img_bytes = 0;
        VO_CUDA(c, cudaMalloc((void**)&s.img_pool, img_total));
        VO_CUDA(c, cudaMalloc((void**)&s.deriv_pool, img_total * sizeof(short2)));
        s.img_bytes = img_total;
        s.pyr.nlevels = 0;
    }
    // the derivative plane must be zero outside the image: clear whenever the geometry changes
    bool same = s.pyr.nlevels == nl;
    for (int l = 0; same && l < nl; l++) same = s.pyr.lv[l].w == ws[l] && s.pyr.lv[l].h == hs[l];
    if (!same) VO_CUDA(c, cudaMemsetAsync(s.deriv_pool, 0, img_total * sizeof(short2), stream));
    s.pyr.nlevels = nl;
    for (int l = 0; l < nl; l++) {
        LevelDesc& d = s.pyr.lv[l];
        d.img = s.img_pool + off[l];
        d.deriv = s.deriv_pool + off[l];
        d.w = ws[l]; d.h = hs[l]; d.stride = st[l];
    }
    if (!same || !s.tmap_valid) encode_level_maps(s);   // (geometry or buffers changed)
    s.pyr.tmaps = s.tmap_valid ? s.tmap_img : nullptr;
    for (int l = 0; l < nl; l++) {
        const LevelDesc& d = s.pyr.lv[l];
        dim3 grid((d.w + 2 * kPad + TX - 1) / TX, (d.h + 2 * kPad + TY - 1) / TY), block(TX, TY);
        if (l == 0)
            VO_CUDA(c, launch_pdl(build_level_kernel<true>, grid, block, stream, (const uint8_t*)s.raw, s.w, d));
        else
            VO_CUDA(c, launch_pdl(build_level_kernel<false>, grid, block, stream, (const uint8_t*)s.pyr.lv[l - 1].img,
                                  s.pyr.lv[l - 1].stride, d));
        c->launches++;
    }
    VO_CUDA(c, cudaGetLastError());
    return VO_OK;
}

}  // namespace vo
