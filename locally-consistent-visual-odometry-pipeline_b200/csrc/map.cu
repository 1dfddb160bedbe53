// trajectory / map bookkeeping of the reference's main loop and a headless
// top-down plot, on the device.
//
// Replaces, per frame, /root/reference/src/main.py:239-255
//     pos = (inv(new_pose) @ [0,0,0,1])[0:3];  poses_plotting = vstack(poses_plotting, pos)
//     X = vision.state['X'];  d = norm(X - pos, axis=1);  X_plotting = vstack(X_plotting, X[(d < 50) & (d > 3)])
// which forces a D2H of all landmarks every frame, and the matplotlib scatter/plot of
// main.py:262-271 / :296-305 (X-Z plane: landmarks blue, travelled path red) by a raster into a
// plain BGR buffer - no window, no graphics API (matplotlib is not on the box either).
//
// The distance test follows numpy bit for bit: d = sqrt((dx*dx + dy*dy) + dz*dz) in float64 without
// FMA contraction.  pos is -R^T t (the pose is rigid), which differs from LAPACK's inverse by rounding
// only (<= 1e-15 relative).
#include "ctx.cuh"

namespace vo {

struct MapScalars {
    int32_t n_traj;     // poses accumulated
    int32_t n_pts;      // landmark rows accumulated
    int32_t last_added; // rows added by the last call
    int32_t overflow;   // rows that did not fit the buffer (host grows it ahead: always 0)
    double pos[3];      // last camera position
    double bounds[4];   // xmin, xmax, zmin, zmax of the last render (with margins)
};

struct MapState {
    DevBuf<double> traj;   // n_traj x 3
    DevBuf<double> pts;    // n_pts x 3
    DevBuf<uint8_t> canvas;
    MapScalars* d_sc = nullptr;
    MapScalars* h_sc = nullptr;   // pinned mirror
    int n_traj_upper = 0, n_pts_upper = 0;
    double* d_T = nullptr;        // staging of a host pose
};

namespace {

// one CTA, order-preserving append of the rows that pass the distance test
__global__ void __launch_bounds__(1024)
map_accumulate_kernel(MapScalars* ms, const DevScalars* sc, const double* __restrict__ Thost,
                      const double* __restrict__ X, int use_prev, int n_upper, double r_min, double r_max,
                      double* __restrict__ traj, int traj_cap, double* __restrict__ pts, int pts_cap)
{
    __shared__ int sWarp[32];
    __shared__ int sBase;
    __shared__ double sPos[3];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const double* T = Thost ? Thost : sc->T;
    if (tid < 3) {
        // pos = -R^T t, accumulated left to right
        const double t0 = T[3], t1 = T[7], t2 = T[11];
        const double v = __dadd_rn(__dadd_rn(__dmul_rn(T[tid], t0), __dmul_rn(T[4 + tid], t1)), __dmul_rn(T[8 + tid], t2));
        sPos[tid] = -v;
    }
    if (tid == 0) sBase = ms->n_pts;
    __syncthreads();
    const int n = min(use_prev ? sc->prev_n_land : sc->n_land, n_upper);
    const double px = sPos[0], py = sPos[1], pz = sPos[2];
    int added = 0;
    for (int base = 0; base < n; base += blockDim.x) {
        const int i = base + tid;
        bool keep = false;
        double x = 0., y = 0., z = 0.;
        if (i < n) {
            x = X[3 * (size_t)i]; y = X[3 * (size_t)i + 1]; z = X[3 * (size_t)i + 2];
            const double dx = __dsub_rn(x, px), dy = __dsub_rn(y, py), dz = __dsub_rn(z, pz);
            const double d = __dsqrt_rn(__dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz)));
            keep = d < r_max && d > r_min;
        }
        const unsigned b = __ballot_sync(0xffffffffu, keep);
        if (lane == 0) sWarp[warp] = __popc(b);
        __syncthreads();
        int off = 0, tot = 0;
        for (int w = 0; w < (int)(blockDim.x >> 5); w++) { if (w < warp) off += sWarp[w]; tot += sWarp[w]; }
        if (keep) {
            const int o = sBase + added + off + __popc(b & ((1u << lane) - 1));
            if (o < pts_cap) { pts[3 * (size_t)o] = x; pts[3 * (size_t)o + 1] = y; pts[3 * (size_t)o + 2] = z; }
        }
        added += tot;
        __syncthreads();
    }
    if (tid == 0) {
        const int k = ms->n_traj;
        if (k < traj_cap) { traj[3 * (size_t)k] = px; traj[3 * (size_t)k + 1] = py; traj[3 * (size_t)k + 2] = pz; ms->n_traj = k + 1; }
        const int fit = min(added, max(pts_cap - sBase, 0));
        ms->n_pts = sBase + fit;
        ms->last_added = fit;
        ms->overflow += added - fit;
        ms->pos[0] = px; ms->pos[1] = py; ms->pos[2] = pz;
    }
}

}  // namespace

// ---- headless top-down plot (main.py:262-271, :296-305) --------------------------------------------
// Raster rules (restated in oracle/map_ref.py, compared pixel for pixel):
//   limits   lo = min - 0.05 (max - min), hi = max + 0.05 (max - min) over landmarks and path
//            (min = max: -/+ 0.5; nothing to draw: [0, 1]) - matplotlib's default 5 % margins
//   pixel    cx = floor(m + (x - xlo) / (xhi - xlo) * (W - 1 - 2m) + 0.5)
//            cy = floor((H - 1 - m) - (z - zlo) / (zhi - zlo) * (H - 1 - 2m) + 0.5)      (z up)
//   canvas   white, black 1 px box around the plot area
//   landmarks  filled discs dx^2 + dy^2 <= r^2, blue;  path: 3 px wide DDA polyline, red, on top
namespace {

__device__ __forceinline__ int fdiv(int a, int b)   // floor division, b > 0
{
    int q = a / b;
    if ((a % b != 0) && (a < 0)) q--;
    return q;
}

__global__ void __launch_bounds__(1024)
map_bounds_kernel(MapScalars* ms, const double* __restrict__ traj, const double* __restrict__ pts)
{
    __shared__ double sm[4][32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nt = ms->n_traj, np_ = ms->n_pts;
    double xmin = INFINITY, xmax = -INFINITY, zmin = INFINITY, zmax = -INFINITY;
    for (int i = tid; i < nt + np_; i += blockDim.x) {
        const double* r = i < nt ? traj + 3 * (size_t)i : pts + 3 * (size_t)(i - nt);
        const double x = r[0], z = r[2];
        xmin = fmin(xmin, x); xmax = fmax(xmax, x); zmin = fmin(zmin, z); zmax = fmax(zmax, z);
    }
    for (int o = 16; o > 0; o >>= 1) {
        xmin = fmin(xmin, __shfl_xor_sync(0xffffffffu, xmin, o)); xmax = fmax(xmax, __shfl_xor_sync(0xffffffffu, xmax, o));
        zmin = fmin(zmin, __shfl_xor_sync(0xffffffffu, zmin, o)); zmax = fmax(zmax, __shfl_xor_sync(0xffffffffu, zmax, o));
    }
    if (lane == 0) { sm[0][warp] = xmin; sm[1][warp] = xmax; sm[2][warp] = zmin; sm[3][warp] = zmax; }
    __syncthreads();
    if (tid == 0) {
        for (int w = 1; w < (int)(blockDim.x >> 5); w++) {
            xmin = fmin(xmin, sm[0][w]); xmax = fmax(xmax, sm[1][w]); zmin = fmin(zmin, sm[2][w]); zmax = fmax(zmax, sm[3][w]);
        }
        double lo[2] = { xmin, zmin }, hi[2] = { xmax, zmax };
        for (int a = 0; a < 2; a++) {
            if (nt + np_ == 0) { lo[a] = 0.; hi[a] = 1.; continue; }
            const double span = __dsub_rn(hi[a], lo[a]);
            if (span == 0.) { lo[a] = __dsub_rn(lo[a], 0.5); hi[a] = __dadd_rn(hi[a], 0.5); }
            else {
                const double mg = __dmul_rn(0.05, span);
                lo[a] = __dsub_rn(lo[a], mg); hi[a] = __dadd_rn(hi[a], mg);
            }
        }
        ms->bounds[0] = lo[0]; ms->bounds[1] = hi[0]; ms->bounds[2] = lo[1]; ms->bounds[3] = hi[1];
    }
}

__device__ __forceinline__ void to_pixel(const MapScalars* ms, double x, double z, int W, int H, int m, int& cx, int& cy)
{
    const double u = __dadd_rn((double)m, __dmul_rn(__ddiv_rn(__dsub_rn(x, ms->bounds[0]), __dsub_rn(ms->bounds[1], ms->bounds[0])),
                                                    (double)(W - 1 - 2 * m)));
    const double v = __dsub_rn((double)(H - 1 - m), __dmul_rn(__ddiv_rn(__dsub_rn(z, ms->bounds[2]), __dsub_rn(ms->bounds[3], ms->bounds[2])),
                                                              (double)(H - 1 - 2 * m)));
    cx = (int)floor(__dadd_rn(u, 0.5));
    cy = (int)floor(__dadd_rn(v, 0.5));
}

__global__ void map_clear_kernel(uint8_t* __restrict__ img, int W, int H, int m)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= W) return;
    const bool in_box = x >= m - 1 && x <= W - m && y >= m - 1 && y <= H - m;
    const bool edge = m >= 1 && in_box && (x == m - 1 || x == W - m || y == m - 1 || y == H - m);
    const uint8_t c = edge ? 0 : 255;
    uint8_t* p = img + 3 * ((size_t)y * W + x);
    p[0] = c; p[1] = c; p[2] = c;
}

__device__ __forceinline__ void put(uint8_t* img, int W, int H, int x, int y, uint8_t b, uint8_t g, uint8_t r)
{
    if (x < 0 || x >= W || y < 0 || y >= H) return;
    uint8_t* p = img + 3 * ((size_t)y * W + x);
    p[0] = b; p[1] = g; p[2] = r;
}

__global__ void map_points_kernel(const MapScalars* ms, const double* __restrict__ pts, uint8_t* __restrict__ img,
                                  int W, int H, int m, int rad)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= ms->n_pts) return;
    int cx, cy;
    to_pixel(ms, pts[3 * (size_t)i], pts[3 * (size_t)i + 2], W, H, m, cx, cy);
    for (int dy = -rad; dy <= rad; dy++)
        for (int dx = -rad; dx <= rad; dx++)
            if (dx * dx + dy * dy <= rad * rad) put(img, W, H, cx + dx, cy + dy, 255, 0, 0);
}

__global__ void map_path_kernel(const MapScalars* ms, const double* __restrict__ traj, uint8_t* __restrict__ img,
                                int W, int H, int m)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    const int nt = ms->n_traj;
    if (k >= max(nt - 1, 1) || nt == 0) return;
    int ax, ay, bx, by;
    to_pixel(ms, traj[3 * (size_t)k], traj[3 * (size_t)k + 2], W, H, m, ax, ay);
    if (nt == 1) { bx = ax; by = ay; }
    else to_pixel(ms, traj[3 * (size_t)(k + 1)], traj[3 * (size_t)(k + 1) + 2], W, H, m, bx, by);
    // clip the walk to something sane if a pose maps far outside (cannot happen: the limits include the path)
    const int dx = bx - ax, dy = by - ay;
    const int n = max(abs(dx), abs(dy));
    for (int i = 0; i <= n; i++) {
        const int x = n ? ax + fdiv(2 * dx * i + n, 2 * n) : ax;
        const int y = n ? ay + fdiv(2 * dy * i + n, 2 * n) : ay;
        for (int oy = -1; oy <= 1; oy++)
            for (int ox = -1; ox <= 1; ox++) put(img, W, H, x + ox, y + oy, 0, 0, 255);
    }
}

}  // namespace

void map_free(Ctx* c)
{
    MapState* m = c->map;
    if (!m) return;
    cudaFree(m->traj.p); cudaFree(m->pts.p); cudaFree(m->canvas.p); cudaFree(m->d_sc); cudaFree(m->d_T);
    if (m->h_sc) cudaFreeHost(m->h_sc);
    delete m;
    c->map = nullptr;
}

static int map_get(Ctx* c, MapState** out)
{
    if (!c->map) {
        MapState* m = new (std::nothrow) MapState();
        if (!m) return set_error(c, VO_ERR_NOMEM, "out of host memory");
        c->map = m;
        VO_CUDA(c, cudaMalloc((void**)&m->d_sc, sizeof(MapScalars)));
        VO_CUDA(c, cudaMemsetAsync(m->d_sc, 0, sizeof(MapScalars), c->stream));
        VO_CUDA(c, cudaMallocHost((void**)&m->h_sc, sizeof(MapScalars)));
        memset(m->h_sc, 0, sizeof(MapScalars));
        VO_CUDA(c, cudaMalloc((void**)&m->d_T, 16 * sizeof(double)));
    }
    *out = c->map;
    return VO_OK;
}

}  // namespace vo

using namespace vo;

extern "C" {

int vo_map_reset(vo_ctx* c)
{
    if (!c || !c->pipe) return VO_ERR_INVALID;
    MapState* m;
    VO_TRY(map_get(c, &m));
    VO_CUDA(c, cudaMemsetAsync(m->d_sc, 0, sizeof(MapScalars), c->stream));
    m->n_traj_upper = 0; m->n_pts_upper = 0;
    return VO_OK;
}

int vo_map_accumulate(vo_ctx* c, const double T44[16], double r_min, double r_max, int use_prev_state, int* n_added)
{
    if (!c || !c->pipe) return VO_ERR_INVALID;
    Pipeline* p = c->pipe;
    MapState* m;
    VO_TRY(map_get(c, &m));
    VO_TRY(join_pose_stream(c));   // the refined pose / promoted rows' 3-D points of a pipelined frame
    int src = p->cur;
    if (use_prev_state) {
        if (p->prev < 0) return set_error(c, VO_ERR_STATE, "no previous state is held");
        src = p->prev;
    }
    // rows of the source buffer: the landmark buffers are sized for the host-side upper bound
    const int n_upper = (int)(p->X[src].cap / 3);
    VO_TRY(ensure(c, m->traj, (size_t)(m->n_traj_upper + 1) * 3, true));
    const int host_bound = use_prev_state ? p->prev_n_upper : p->n_land_upper;
    const int bound = host_bound < n_upper ? host_bound : n_upper;
    VO_TRY(ensure(c, m->pts, (size_t)(m->n_pts_upper + bound + 1) * 3, true));
    const double* dT = nullptr;
    if (T44) {
        VO_CUDA(c, cudaMemcpyAsync(m->d_T, T44, 16 * sizeof(double), cudaMemcpyDefault, c->stream));
        dT = m->d_T;
    }
    map_accumulate_kernel<<<1, 1024, 0, c->stream>>>(m->d_sc, p->d_sc, dT, p->X[src].p, use_prev_state ? 1 : 0, bound, r_min, r_max,
                                                     m->traj.p, (int)(m->traj.cap / 3), m->pts.p, (int)(m->pts.cap / 3));
    c->launches++;
    VO_CUDA(c, cudaGetLastError());
    m->n_traj_upper += 1;
    m->n_pts_upper += bound;
    if (n_added || T44) {   // a pageable host pose must have been consumed before returning
        VO_CUDA(c, cudaMemcpyAsync(m->h_sc, m->d_sc, sizeof(MapScalars), cudaMemcpyDeviceToHost, c->stream));
        VO_CUDA(c, cudaStreamSynchronize(c->stream));
        m->n_pts_upper = m->h_sc->n_pts;   // tighten
        if (n_added) *n_added = m->h_sc->last_added;
    }
    return VO_OK;
}

int vo_map_get(vo_ctx* c, double* traj, int traj_cap, int* n_traj, double* pts, int pts_cap, int* n_pts, double pos[3])
{
    if (!c || !c->pipe) return VO_ERR_INVALID;
    MapState* m;
    VO_TRY(map_get(c, &m));
    VO_CUDA(c, cudaMemcpyAsync(m->h_sc, m->d_sc, sizeof(MapScalars), cudaMemcpyDeviceToHost, c->stream));
    VO_CUDA(c, cudaStreamSynchronize(c->stream));
    const int nt = m->h_sc->n_traj, np_ = m->h_sc->n_pts;
    m->n_pts_upper = np_;
    if (n_traj) *n_traj = nt;
    if (n_pts) *n_pts = np_;
    if (pos) memcpy(pos, m->h_sc->pos, 3 * sizeof(double));
    if ((traj && nt > traj_cap) || (pts && np_ > pts_cap))
        return set_error(c, VO_ERR_CAPACITY, "map has %d poses / %d landmarks, buffers %d / %d", nt, np_, traj_cap, pts_cap);
    if (traj && nt > 0) VO_CUDA(c, cudaMemcpyAsync(traj, m->traj.p, sizeof(double) * 3 * nt, cudaMemcpyDefault, c->stream));
    if (pts && np_ > 0) VO_CUDA(c, cudaMemcpyAsync(pts, m->pts.p, sizeof(double) * 3 * np_, cudaMemcpyDefault, c->stream));
    VO_CUDA(c, cudaStreamSynchronize(c->stream));
    return VO_OK;
}

int vo_map_render(vo_ctx* c, uint8_t* bgr, int w, int h, int margin, int radius, double bounds[4])
{
    if (!c || !c->pipe || !bgr) return VO_ERR_INVALID;
    if (w < 8 || h < 8 || w > 8192 || h > 8192 || margin < 0 || 2 * margin >= (w < h ? w : h) - 1 || radius < 0 || radius > 64)
        return set_error(c, VO_ERR_INVALID, "bad canvas %dx%d margin %d radius %d", w, h, margin, radius);
    MapState* m;
    VO_TRY(map_get(c, &m));
    VO_TRY(ensure(c, m->canvas, (size_t)w * h * 3));
    VO_TRY(ensure(c, m->traj, 3)); VO_TRY(ensure(c, m->pts, 3));
    map_bounds_kernel<<<1, 1024, 0, c->stream>>>(m->d_sc, m->traj.p, m->pts.p);
    map_clear_kernel<<<dim3((w + 255) / 256, h), 256, 0, c->stream>>>(m->canvas.p, w, h, margin);
    const int np_ = m->n_pts_upper > 0 ? m->n_pts_upper : 1, nt = m->n_traj_upper > 0 ? m->n_traj_upper : 1;
    map_points_kernel<<<(np_ + 255) / 256, 256, 0, c->stream>>>(m->d_sc, m->pts.p, m->canvas.p, w, h, margin, radius);
    map_path_kernel<<<(nt + 127) / 128, 128, 0, c->stream>>>(m->d_sc, m->traj.p, m->canvas.p, w, h, margin);
    c->launches += 4;
    VO_CUDA(c, cudaGetLastError());
    VO_CUDA(c, cudaMemcpyAsync(bgr, m->canvas.p, (size_t)w * h * 3, cudaMemcpyDefault, c->stream));
    VO_CUDA(c, cudaMemcpyAsync(m->h_sc, m->d_sc, sizeof(MapScalars), cudaMemcpyDeviceToHost, c->stream));
    VO_CUDA(c, cudaStreamSynchronize(c->stream));
    if (bounds) memcpy(bounds, m->h_sc->bounds, 4 * sizeof(double));
    return VO_OK;
}

}  // extern "C"
