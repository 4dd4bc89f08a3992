// Per-pose kernels: one CTA renders (and, in COST mode, scores) one candidate pose.
//
// MODE_COST  -- fused render -> cost: the COST stage of render_cuda_multi_unified (renderer.cu:1431-1934) without
//               materialising rendered images, clouds, KNN arrays or the P x N_obs `explained` matrix in HBM.
// MODE_IMAGE -- the rasterizer alone (image_render, image_renderer.cuh:336-496): writes int32 depth + RGB images,
//               either full resolution (stage RENDER/DEBUG) or only the stride lattice (stage CLOUD).
//
// Steps of one CTA:
//   1. tile:    conservative screen bbox of the posed model AABB -> rectangle of the stride lattice
//   2. raster:  every triangle of the pose's model is transformed with the reference's exact float sequence
//               (exact_math.cuh); only pixels ON the stride lattice can ever reach the cost (compute_point_clouds.cuh
//               :47-56 samples x*stride,y*stride), so only those are z-tested, with atomicMin into a shared-memory
//               lattice z-buffer.  Triangles that cover no lattice pixel (the vast majority for fine meshes) are
//               culled hierarchically and conservatively (cluster bounding sphere, then an approximate per-triangle
//               bbox); survivors are compacted into a shared queue and rasterised exactly by dense warps.
//   3. cost:    per lattice pixel: source-occlusion rule, back-projection, exact nearest observed point through the
//               uniform grid, explained-bit set in a shared bitmask; warp-shuffle reduction of the three counters;
//               the reference's percentage formulae (compute_costs.cuh:364-378,407-446).
#include <cuda_runtime.h>

#include "clusters.hpp"
#include "context.hpp"
#include "kernels.cuh"

namespace perch {

namespace {

#ifndef PERCH_POSE_THREADS
#define PERCH_POSE_THREADS 128
#endif
constexpr int kThreads = PERCH_POSE_THREADS;
constexpr int kWarps = kThreads / 32;
#ifndef PERCH_POSE_MIN_CTAS
#define PERCH_POSE_MIN_CTAS (1024 / PERCH_POSE_THREADS)
#endif
constexpr int kMinCtas = PERCH_POSE_MIN_CTAS;  // resident CTAs per SM the register budget is sized for (32 warps)
constexpr int kQueueCap = 16 * kThreads;   // candidate triangles buffered between exact passes
#ifndef PERCH_LANE_QUADS
#define PERCH_LANE_QUADS 4
#endif
constexpr int kLaneQuads = PERCH_LANE_QUADS;  // quads sphere-tested per lane per claim (independent loads in flight)
constexpr int kClaim = 32 * kLaneQuads;  // quads per warp claim
constexpr int kPendCap = kClaim + 32;    // per-warp list of surviving quads awaiting the per-triangle test (< 32 carried)
constexpr int kClaimMaxTris = kPendCap * 2;  // most candidates one claim can append (incl. carried quads)
static_assert(kClusterTris == 4 && kClusterQuads == 2, "quad q = 2 * cluster + (0 | 1), triangle slot = 2 * q + (0 | 1)");
constexpr int kBigCells = 48;     // lattice pixels above which a triangle is rasterised by the whole CTA
constexpr int kBigCap = 64;
// Bounds on |approximate - exact| screen coordinates (pixels) assumed by the culling levels.  tools/approx_error.cu
// measures the actual difference: <= 1.3e-3 px over 2e8 in-view vertices of objects 0.2-3 m away, <= 1.9e-4 px beyond;
// it grows like 1/z for vertices close to the camera plane, so a vertex with |z| <= kNearZ (pose units, cm) sends its
// triangle to the exact path untested, and level 3 scales its bound by max(1, kRefZ / |z|).
constexpr float kEps = 0.02f;     // levels 1 and 2 (bbox tests)
constexpr float kEps3 = 0.006f;   // level 3 (edge-function test)
constexpr float kNearZ = 1.0f;
constexpr float kRefZ = 5.0f;

__device__ __forceinline__ float __frcp_approx(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

// atomicAdd on a shared counter from ONE elected lane.  The compiler cannot see that a single lane is active and wraps
// every atomicAdd with a uniform address in a vote / popc / shuffle aggregation (13 instructions, ncu r2_v14: 4.7 % of the
// kernel's warp-instructions); the plain instruction is all that is needed here.
__device__ __forceinline__ int smem_add_elected(int* p, int v) {
    int old;
    unsigned lane_hi;  // always 0; keeps ptxas from proving the address warp-uniform (it aggregates inline PTX too)
    asm volatile("{ .reg .u32 l; mov.u32 l, %%laneid; shr.u32 %0, l, 5; }" : "=r"(lane_hi));
    asm volatile("atom.shared.add.s32 %0, [%1], %2;" : "=r"(old) : "r"((unsigned)__cvta_generic_to_shared(p) + lane_hi), "r"(v) : "memory");
    return old;
}

// `make check` build (tools/bounds_check.py): every data-dependent index into the kernel's shared-memory arrays is range
// checked; a violation sets bit 8 of PoseArgs::err_flag (reported as PERCH_ERR_CUDA) and the access is skipped.  Stands in
// for compute-sanitizer memcheck, which is closed on this pool.  The product build compiles the checks away.
#ifdef PERCH_BOUNDS_CHECK
#define IDX_OK(i, cap) (((i) >= 0 && (i) < (cap)) ? true : (atomicOr(a.err_flag, 256), false))
#else
#define IDX_OK(i, cap) true
#endif

#ifdef PERCH_STATS
// Counters build only (`make stats`, tools/pose_stats.py): survivors of each culling level, summed over all poses.
__device__ unsigned long long g_stats[16];
#define STAT(i, v) atomicAdd(&g_stats[i], (unsigned long long)(v))
#else
#define STAT(i, v) ((void)0)
#endif

template <bool COLOR> struct ZCell;
template <> struct ZCell<false> {
    using type = unsigned int;
    static __device__ __forceinline__ type empty() { return 0xFFFFFFFFu; }
    // order-preserving map int32 -> uint32 so that atomicMin on the key is a signed min on depth
    static __device__ __forceinline__ type make(int32_t d, uint32_t) { return (uint32_t)d ^ 0x80000000u; }
    static __device__ __forceinline__ int32_t depth(type k) { return (int32_t)(k ^ 0x80000000u); }
    static __device__ __forceinline__ uint32_t tri(type) { return 0; }
};
template <> struct ZCell<true> {
    using type = unsigned long long;
    static __device__ __forceinline__ type empty() { return ~0ull; }
    // (depth, triangle index): the lowest-index triangle among equal depths wins the colour (canonical tie-break)
    static __device__ __forceinline__ type make(int32_t d, uint32_t t) {
        return ((unsigned long long)((uint32_t)d ^ 0x80000000u) << 32) | t;
    }
    static __device__ __forceinline__ int32_t depth(type k) { return (int32_t)((uint32_t)(k >> 32) ^ 0x80000000u); }
    static __device__ __forceinline__ uint32_t tri(type k) { return (uint32_t)k; }
};

struct LatticeRect {
    int x0, x1, y0, y1;  // inclusive lattice coordinates; empty when x0 > x1 or y0 > y1
};

// Lattice pixels visited by the reference's bbox loops (image_renderer.cuh:110-111) for this triangle.
// floor(a / s) for 0 <= a < 2^23 without the integer-division sequence: float estimate, exact correction
__device__ __forceinline__ int div_floor(int a, int s, float inv_s) {
    int q = (int)((float)a * inv_s);
    const int r = a - q * s;
    if (r < 0) --q;
    else if (r >= s) ++q;
    return q;
}

__device__ __forceinline__ LatticeRect lattice_rect_of(float bminx, float bmaxx, float bminy, float bmaxy, int W, int H,
                                                       int s, float inv_s) {
    LatticeRect r{1, 0, 1, 0};
    if (!(bmaxx >= 0.f) || !(bmaxy >= 0.f)) return r;  // also rejects NaN upper bounds (loop condition false)
    unsigned long long p0s = (unsigned long long)__fadd_rn(bminx, 0.5f);  // size_t(bboxmin+0.5f); NaN -> 0
    unsigned long long p1s = (unsigned long long)__fadd_rn(bminy, 0.5f);
    if (p0s > (unsigned long long)(W - 1) || p1s > (unsigned long long)(H - 1)) return r;
    int p0e = (int)bmaxx, p1e = (int)bmaxy;  // <= W-1 / H-1 by the clamp
    if ((int)p0s > p0e || (int)p1s > p1e) return r;
    r.x0 = div_floor((int)p0s + s - 1, s, inv_s);
    r.x1 = div_floor(p0e, s, inv_s);
    // raster row P1 is image row H-1-P1 (image_renderer.cuh:128)
    r.y0 = div_floor(H - 1 - p1e + s - 1, s, inv_s);
    r.y1 = div_floor(H - 1 - (int)p1s, s, inv_s);
    return r;
}

// Candidate entry = triangle slot t (4 * cluster + j: triangle j & 1 of quad j >> 1, clusters.hpp) in the low kSlotBits
// bits and, above them, the slot's 6-bit permutation: which three of the quad's four vertices the triangle uses, in its
// ORIGINAL order.  Carrying the permutation spares the exact path a dependent load.
constexpr int kSlotBits = 26;
constexpr int kSlotMask = (1 << kSlotBits) - 1;
__device__ __forceinline__ void load_tri(const ModelsDev& m, int entry, float v[9]) {
    const int t = entry & kSlotMask;
    const uint32_t perm = (uint32_t)entry >> kSlotBits;
    // nine 4-byte loads at perm-dependent offsets.  (Three 16-byte loads of the whole quad + selects measured slower:
    // the exact path is issue-bound and the 27 selects cost more than the loads save: C3 1.93 -> 1.71 ms, C2 1.585 ->
    // 1.509 ms with this form; the level-2 test, which uses all twelve floats, keeps the 16-byte loads.)
    const float* base = m.cl_recs + (int64_t)(t >> 1) * kQuadFloats;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float* p = base + 3 * ((perm >> (2 * k)) & 3u);
        v[3 * k] = __ldg(p); v[3 * k + 1] = __ldg(p + 1); v[3 * k + 2] = __ldg(p + 2);
    }
}

// CIEDE2000 between two CIELAB triples with the reference's mixed float/double evaluation
// (compute_costs.cuh:90-159): float libm calls on double operands, and the two hue branches that are NOT
// mutually exclusive with the low-chroma case (the `if` at :112/:133 is overwritten by the following if/else).
__device__ double ciede2000_ref(float L1, float a1, float b1, float L2, float a2, float b2) {
    const double kPi = 3.14159265358979323846;
    double c1 = sqrtf(a1 * a1 + b1 * b1);
    double c2 = sqrtf(a2 * a2 + b2 * b2);
    double mc = (c1 + c2) / 2.0;
    double mc7 = (mc * mc * mc) * (mc * mc * mc) * mc;
    double g = 0.5 * (1 - sqrtf(mc7 / (mc7 + 6103515625.)));
    double a1p = a1 * (1 + g);
    double a2p = a2 * (1 + g);
    c1 = sqrtf(a1p * a1p + b1 * b1);
    c2 = sqrtf(a2p * a2p + b2 * b2);
    double h1 = fmodf(atan2f(b1, a1p) + 2 * kPi, 2 * kPi);
    double h2 = fmodf(atan2f(b2, a2p) + 2 * kPi, 2 * kPi);
    double dL = L2 - L1;
    double dC = c2 - c1;
    double dh = h2 - h1;
    if (fabs(dh) > kPi) dh = (h2 > h1) ? dh - 2 * kPi : dh + 2 * kPi;
    double dH = 2 * sqrtf(c1 * c2) * sinf(dh / 2);
    double mL = (L1 + L2) / 2;
    mc = (c1 + c2) / 2.0;
    mc7 = (mc * mc * mc) * (mc * mc * mc) * mc;
    double mH;
    if (fabs(h1 - h2) <= kPi + 1e-5) mH = (h1 + h2) / 2;
    else if (h1 + h2 < 2 * kPi) mH = (h1 + h2 + 2 * kPi) / 2;
    else mH = (h1 + h2 - 2 * kPi) / 2;
    double T = 1 - 0.17 * cosf(mH - (30) * kPi / 180) + 0.24 * cosf(2 * mH) + 0.32 * cosf(3 * mH + (6) * kPi / 180) -
               0.2 * cosf(4 * mH - (63) * kPi / 180);
    double dl50 = (mL - 50) * (mL - 50);
    double sl = 1 + (0.015 * dl50) / sqrtf(20 + dl50);
    double sc = 1 + 0.045 * mc;
    double sh = 1 + 0.015 * mc * T;
    double rc = 2 * sqrtf(mc7 / (mc7 + 6103515625.));
    double q = ((mH) / kPi * 180 - 275) / 25;
    double rt = -sinf((60 * expf(-(q * q))) * kPi / 180) * rc;
    double tl = dL / sl, tc = dC / sc, th = dH / sh;
    return sqrtf(tl * tl + tc * tc + th * th + rt * dC / sc * dH / sh);
}

// Exact nearest observed point among those that can lie within sqrt(res2) of q.  Returns the squared distance of the
// best candidate seen (FLT_MAX if none) and its reference index; lowest reference index wins ties, which is what a
// strict-`<` linear scan (the brute force this replaces) yields.
__device__ __forceinline__ void nn_query(const ObsGridDev& g, float qx, float qy, float qz, bool use_label, int label,
                                         float& best, int& best_ref) {
    best = FLT_MAX;
    best_ref = -1;
    float fx = floorf(__fmul_rn(__fsub_rn(qx, g.ox), g.inv_h));
    float fy = floorf(__fmul_rn(__fsub_rn(qy, g.oy), g.inv_h));
    float fz = floorf(__fmul_rn(__fsub_rn(qz, g.oz), g.inv_h));
    if (!(fx >= -1.f && fx <= (float)g.dx && fy >= -1.f && fy <= (float)g.dy && fz >= -1.f && fz <= (float)g.dz)) return;
    int cx = (int)fx, cy = (int)fy, cz = (int)fz;
    int x0 = max(cx - 1, 0), x1 = min(cx + 1, g.dx - 1);
    if (x0 > x1) return;
    const int y0 = max(cy - 1, 0), y1 = min(cy + 1, g.dy - 1);
    for (int z = max(cz - 1, 0); z <= min(cz + 1, g.dz - 1); ++z) {
        // the (up to) three x-runs of this z-slice: all six range bounds are requested before any point is read
        uint32_t rb[3], re[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const int y = y0 + k;
            const int row = (z * g.dy + min(y, y1)) * g.dx;
            rb[k] = __ldg(g.cell_start + row + x0);
            re[k] = (y <= y1) ? __ldg(g.cell_start + row + x1 + 1) : rb[k];
        }
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            for (uint32_t j = rb[k]; j < re[k]; ++j) {
                // label and point are requested together (the label test does not gate the point's load)
                const int pl = use_label ? __ldg(g.label + j) : label;
                const float4 p = __ldg(g.pts + j);
                if (pl != label) continue;
                float d = sqdist3(qx, qy, qz, p.x, p.y, p.z);
                int ref = __float_as_int(p.w);
                if (d < best || (d == best && ref < best_ref)) {
                    best = d;
                    best_ref = ref;
                }
            }
        }
    }
}

enum { MODE_COST = 0, MODE_IMAGE = 1 };

template <int MODE, bool COLOR>
__global__ void __launch_bounds__(kThreads, kMinCtas) pose_kernel(PoseArgs a) {
    using Z = ZCell<COLOR>;
    using zkey = typename Z::type;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    // layout: [candidate queue][per-warp pending quads] at compile-time offsets (the hot loops of pass A address them
    // without pointer arithmetic), then the z-buffer (tile_cap cells) and the explained-point bitmask (run-time sizes)
    int* queue = reinterpret_cast<int*>(smem_raw);
    int* wpend = queue + kQueueCap + (threadIdx.x >> 5) * kPendCap;  // per-warp quads that survived level 1
    constexpr size_t kFixedBytes = (size_t)(kQueueCap + kWarps * kPendCap) * sizeof(int);
    static_assert(kFixedBytes % 16 == 0, "the z-buffer behind the fixed part must stay 16-byte aligned");
    zkey* zbuf = reinterpret_cast<zkey*>(smem_raw + kFixedBytes);
    const int mask_words = (MODE == MODE_COST) ? a.mask_words : 0;
    uint32_t* mask = reinterpret_cast<uint32_t*>(smem_raw + kFixedBytes + (size_t)a.tile_cap * sizeof(zkey));

    __shared__ Mat4 s_pose;
    __shared__ int s_rect[4];
    __shared__ int s_qcount, s_bigcount, s_next;
    __shared__ float s_apx[20];
    __shared__ int s_big[kBigCap];
    __shared__ int s_n, s_c, s_e;
    __shared__ float s_band[4];
    __shared__ float s_minz;  // smallest camera-space depth of the posed model AABB (0: unknown / not in front)

    const int pose_i = blockIdx.x;
    const int tid = threadIdx.x;
    const int W = a.cam.width, H = a.cam.height, s = a.stride;
    const int Lw = W / s, Lh = (H + s - 1) / s;
    const float wf = (float)W, hf = (float)H;  // cvt.rn.f32.u64 of width/height
    const float half_w = 0.5f * wf, half_h = 0.5f * hf, inv_s = 1.0f / (float)s;
    // lattice raster rows are P1 = H-1-k*s == row_phase (mod s) with row_phase in [0, s)
    const float row_phase = (float)((H - 1) % s);
    const float cmaxx = (float)(W - 1), cmaxy = (float)(H - 1);

    if (tid < 16) reinterpret_cast<float*>(&s_pose)[tid] = reinterpret_cast<const float*>(a.poses + pose_i)[tid];
    if (tid == 0) { s_n = 0; s_c = 0; s_e = 0; s_qcount = 0; s_bigcount = 0; }
    for (int w = tid; w < mask_words; w += kThreads) mask[w] = 0u;
    __syncthreads();

    const int model = a.pose_model_map[pose_i];
    const bool model_ok = model >= 0 && model < a.models.n_models;
    const int t_low = model_ok ? a.models.model_low[model] : 0;
    const int t_high = model_ok ? a.models.model_high[model] : 0;
    const int cl_low = model_ok ? a.models.cl_low[model] : 0;
    const int cl_high = model_ok ? a.models.cl_high[model] : 0;
    const int lane = tid & 31;
    const bool use_label = a.pose_label != nullptr;
    const int pose_label = use_label ? a.pose_label[pose_i] : 0;

    // ---- 1. lattice tile from the posed AABB (first warp) -------------------------------------------------
    if (tid < 32) {
        float minx = FLT_MAX, maxx = -FLT_MAX, miny = FLT_MAX, maxy = -FLT_MAX, minz = FLT_MAX;
        bool bad = false;
        if (tid < 8 && model_ok) {
            const float* bb = a.models.model_aabb + 6 * model;
            float x = bb[(tid & 1) ? 3 : 0], y = bb[(tid & 2) ? 4 : 1], z = bb[(tid & 4) ? 5 : 2];
            ScreenVertex v = transform_vertex(s_pose, a.cam.proj, wf, hf, x, y, z);
            bad = !(v.z > 1e-6f) || !(fabsf(v.sx) < 1e7f) || !(fabsf(v.sy) < 1e7f);
            minx = maxx = v.sx; miny = maxy = v.sy; minz = v.z;
        }
#pragma unroll
        for (int o = 4; o >= 1; o >>= 1) {
            minx = fminf(minx, __shfl_xor_sync(0xffffffffu, minx, o));
            maxx = fmaxf(maxx, __shfl_xor_sync(0xffffffffu, maxx, o));
            miny = fminf(miny, __shfl_xor_sync(0xffffffffu, miny, o));
            maxy = fmaxf(maxy, __shfl_xor_sync(0xffffffffu, maxy, o));
            minz = fminf(minz, __shfl_xor_sync(0xffffffffu, minz, o));
        }
        bad = __any_sync(0xffffffffu, bad);
        if (tid == 0) {
            s_minz = (bad || !model_ok) ? 0.f : minz;
            int x0 = 0, x1 = Lw - 1, y0 = 0, y1 = Lh - 1;
            if (!model_ok || t_high <= t_low || cl_high <= cl_low) { x0 = 1; x1 = 0; }
            else if (!bad) {
                // 2 px guard band against the float rounding of the per-vertex path; raster row -> image row flip
                float px0 = floorf(minx) - 2.f, px1 = ceilf(maxx) + 2.f;
                float py0 = floorf(miny) - 2.f, py1 = ceilf(maxy) + 2.f;
                int ix0 = (int)fmaxf(px0, 0.f), ix1 = (int)fminf(px1, cmaxx);
                int r0 = (int)fmaxf(py0, 0.f), r1 = (int)fminf(py1, cmaxy);
                if (px1 < 0.f || px0 > cmaxx || py1 < 0.f || py0 > cmaxy) { x0 = 1; x1 = 0; }
                else {
                    x0 = (ix0 + s - 1) / s; x1 = ix1 / s;
                    y0 = (H - 1 - r1 + s - 1) / s; y1 = (H - 1 - r0) / s;
                    x1 = min(x1, Lw - 1); y1 = min(y1, Lh - 1);
                }
            }
            s_rect[0] = x0; s_rect[1] = x1; s_rect[2] = y0; s_rect[3] = y1;
        }
    }
    if (tid == 32) {
        const Mat4& pose = s_pose;
        // Per-pose constants of the APPROXIMATE transform used by the conservative culling (pass A): lattice
        // coordinates (ux, uy) = (sx / s, (sy - row_phase) / s), so that lattice pixels are the integer points, as
        //   ux = (U . (x,y,z,1)) / (C . (x,y,z,1)) + offx,   uy = (V . (x,y,z,1)) / (C . (x,y,z,1)) + offy
        // with U = hw/s * (P.a0*pose.a + P.a1*pose.b) (+ P.a3 on the translation), offx = hw*(1 + P.a2)/s, C = pose.c
        // (the exact path, exact_math.cuh, evaluates sx = fma(W, 0.5, ((P.a . cam) / cam.z * W) * 0.5) instead).
        const Mat4& P = a.cam.proj;
        const float kx = half_w * inv_s, ky = half_h * inv_s;
        float* U = s_apx, *V = s_apx + 4, *C = s_apx + 8;
        U[0] = kx * (P.a0 * pose.a0 + P.a1 * pose.b0); U[1] = kx * (P.a0 * pose.a1 + P.a1 * pose.b1);
        U[2] = kx * (P.a0 * pose.a2 + P.a1 * pose.b2); U[3] = kx * (P.a0 * pose.a3 + P.a1 * pose.b3 + P.a3);
        V[0] = ky * (P.b0 * pose.a0 + P.b1 * pose.b0); V[1] = ky * (P.b0 * pose.a1 + P.b1 * pose.b1);
        V[2] = ky * (P.b0 * pose.a2 + P.b1 * pose.b2); V[3] = ky * (P.b0 * pose.a3 + P.b1 * pose.b3 + P.b3);
        C[0] = pose.c0; C[1] = pose.c1; C[2] = pose.c2; C[3] = pose.c3;
        s_apx[12] = (half_w + half_w * P.a2) * inv_s;
        s_apx[13] = (half_h + half_h * P.b2 - row_phase) * inv_s;
        // cluster test: a point within r of a centre with numerators (N, Z) moves ux by at most
        //   r * (|U| + |N/Z| * |C|) / (Z - r*|C|)        (|.| = norm of the 3-vector part)
        s_apx[14] = sqrtf(U[0] * U[0] + U[1] * U[1] + U[2] * U[2]) * 1.0001f;
        s_apx[15] = sqrtf(V[0] * V[0] + V[1] * V[1] + V[2] * V[2]) * 1.0001f;
        s_apx[16] = sqrtf(C[0] * C[0] + C[1] * C[1] + C[2] * C[2]) * 1.0001f;
    }
    __syncthreads();
    const int rx0 = s_rect[0], rx1 = s_rect[1], ry0 = s_rect[2], ry1 = s_rect[3];
    const int tile_w = rx1 - rx0 + 1;
    const bool tile_empty = (rx0 > rx1) || (ry0 > ry1);
    int n_local = 0, c_local = 0;
    int tree_flags_local = 0, tree_clutter_local = 0, tree_total_local = 0;  // MODE_IMAGE with a.tree_total

    if (!tile_empty) {
        const int band_rows = max(1, a.tile_cap / tile_w);
        for (int by0 = ry0; by0 <= ry1; by0 += band_rows) {
            const int by1 = min(by0 + band_rows - 1, ry1);
            const int band_cells = tile_w * (by1 - by0 + 1);
            for (int c = tid; c < band_cells; c += kThreads) zbuf[c] = Z::empty();
            if (tid == 0) {
                s_qcount = 0; s_bigcount = 0;
                // Band limits in lattice coordinates for the candidate test: columns ux in [rx0, rx1]; raster rows
                // P1 = H-1-ly*s for ly in [by0, by1], i.e. uy = (H-1-row_phase)/s - ly.  Computed once and handed to every
                // thread through shared memory: as expressions of (rx0, by0, inv_s ...) the compiler re-derived them inside
                // the culling loops (I2FP + FFMA per use).  C2 1.245 -> 1.238 ms; doing the same with the margins and the
                // sphere constants (m_lo3, Kg ...) was measured too and is slower again (1.249 ms).
                const float rows_k = (float)((H - 1) / s);
                s_band[0] = (float)rx0 - 0.5f * inv_s; s_band[1] = (float)rx1;
                s_band[2] = (rows_k - (float)by1) - 0.5f * inv_s; s_band[3] = rows_k - (float)by0;
            }
            __syncthreads();
            const float band_lo_x = s_band[0], band_hi_x = s_band[1], band_lo_y = s_band[2], band_hi_y = s_band[3];

            // Exact path for one triangle: the reference's float sequence (exact_math.cuh), its bbox loops restricted
            // to lattice pixels of this band, z-test with atomicMin.  `coop` = all threads of the CTA share the work.
            // ONE copy of it exists in the kernel (`coop` is a run-time value): the kernel is 100 KB of SASS and stalls
            // on instruction fetch, so the per-thread and the cooperative pass share the code.
            auto raster_tri = [&](int entry, bool coop) {
                float v[9];
                load_tri(a.models, entry, v);
                ScreenVertex A = transform_vertex(s_pose, a.cam.proj, wf, hf, v[0], v[1], v[2]);
                ScreenVertex B = transform_vertex(s_pose, a.cam.proj, wf, hf, v[3], v[4], v[5]);
                ScreenVertex C = transform_vertex(s_pose, a.cam.proj, wf, hf, v[6], v[7], v[8]);
                float bminx, bmaxx, bminy, bmaxy;
                bbox_of(A, B, C, cmaxx, cmaxy, bminx, bmaxx, bminy, bmaxy);
                LatticeRect r = lattice_rect_of(bminx, bmaxx, bminy, bmaxy, W, H, s, inv_s);
                if (r.x0 > r.x1 || r.y0 > r.y1) return;
                if (r.x0 < rx0 || r.x1 > rx1 || r.y0 < ry0 || r.y1 > ry1) atomicOr(a.err_flag, 1);
                int x0 = max(r.x0, rx0), x1 = min(r.x1, rx1), y0 = max(r.y0, by0), y1 = min(r.y1, by1);
                if (x0 > x1 || y0 > y1) return;
                const int w = x1 - x0 + 1, cells = w * (y1 - y0 + 1);
                if (!coop && cells > kBigCells) {
                    // many lattice pixels: left to the cooperative pass (unless its list is full)
                    const int bslot = atomicAdd(&s_bigcount, 1);
                    if (bslot < kBigCap) {
                        if (IDX_OK(bslot, kBigCap)) s_big[bslot] = entry;
                        return;
                    }
                }
                if (!coop || tid == 0) { STAT(4, 1); STAT(5, cells); if (coop) STAT(7, 1); }
                TriSetup ts;
                setup_edges(A, B, C, ts);
                // original triangle index (colour tie-break and lookup), or the pose index for perch_compose_source
                const uint32_t id = !COLOR ? 0u : (MODE == MODE_IMAGE && a.key_pose_id) ? (uint32_t)pose_i
                                                : (uint32_t)__ldg(a.models.cl_orig + (entry & kSlotMask));
                // lattice pixels c = first, first + step, ... of the clipped rectangle (row-major); a lone thread takes
                // every pixel, the CTA every kThreads-th.  Division only where the step is not 1.
                int lx = x0, ly = y0, c = 0, step_rows = 0, step_cols = 1, step = 1;
                if (coop) {
                    c = tid; step = kThreads;
                    ly = y0 + c / w; lx = x0 + c % w;
                    step_rows = kThreads / w; step_cols = kThreads % w;
                }
                for (; c < cells; c += step) {
                    int32_t d;
                    if (fragment_depth(ts, (float)(lx * s), (float)(H - 1 - ly * s), d)) {
                        if (IDX_OK((ly - by0) * tile_w + (lx - rx0), a.tile_cap)) atomicMin(&zbuf[(ly - by0) * tile_w + (lx - rx0)], Z::make(d, id));
                        STAT(6, 1);
                    }
                    lx += step_cols; ly += step_rows;
                    if (lx > x1) { lx -= w; ++ly; }
                }
            };

            // Conservative culling (pass A), in lattice coordinates.  The question at every level: can the reference's pixel loops,
            // P in (min-0.5, max] clamped to the image (image_renderer.cuh:110-111), visit a lattice pixel of this band,
            // when the true bounds may differ from (minx..maxy) by kEps pixels?  I.e. is there an integer in
            // [min - m_lo - 1e-3, max + m_hi] along both axes, inside the band.
            const float m_lo = (0.5f + kEps) * inv_s, m_hi = kEps * inv_s;
            const float m_lo3 = m_lo + 1e-3f, clamp_lo_x = band_lo_x - 1e-3f, clamp_lo_y = band_lo_y - 1e-3f;
            const float U0 = s_apx[0], U1 = s_apx[1], U2 = s_apx[2], U3 = s_apx[3];
            const float V0 = s_apx[4], V1 = s_apx[5], V2 = s_apx[6], V3 = s_apx[7];
            const float C0 = s_apx[8], C1 = s_apx[9], C2 = s_apx[10], C3 = s_apx[11];
            const float offx = s_apx[12], offy = s_apx[13];
            // Level 1: bounding sphere of a QUAD (two edge-sharing triangles; clusters.cpp).  A point within r of a centre c
            // with camera depth Z_c moves the lattice coordinate ux by at most r (|U| + |ux_c| |C|) / (Z_c - r |C|)   (|.| =
            // norm of the 3-vector part; ux_c relative to the principal point).  Per pose: centres lie inside the model AABB,
            // so Z_c >= zmin, the nearest AABB corner; with rmax the model's largest quad radius,
            // 1 / (Z_c - r |C|) <= g / Z_c for ONE per-pose g = 1 / (1 - rmax |C| / zmin): the second reciprocal of the
            // textbook form disappears and the projected radius is (r / Z_c) (|U| g + |ux_c| |C| g).  When the AABB is not
            // safely in front of the camera (zmin <= 4 rmax |C|) g = inf: every quad passes to level 2, which sends
            // near-plane vertices to the exact path.
            // NaN / inf anywhere makes the rejecting comparisons false: kept.  (Round 1 tested the spheres of 4-triangle
            // clusters here: at ~5 px against the 8 px lattice 52 % survived; a quad sphere is ~3.9 px and 30 % survive.)
            float Kg, KUg, KVg;  // |C| g, |U| g, |V| g
            {
                const float nC = s_apx[16], zmin = s_minz;
                const float rmax = model_ok ? __ldg(a.models.model_qrmax + model) : 0.f;
                const bool safe = zmin > 4.f * rmax * nC;
                const float g = safe ? 1.001f / (1.f - rmax * nC / fmaxf(zmin, 1e-30f)) : INFINITY;
                Kg = nC * g; KUg = s_apx[14] * g; KVg = s_apx[15] * g;
            }
            const float offx_lo = offx - m_lo3, offy_lo = offy - m_lo3;
            auto sphere_rejects = [&](const float4 b) -> bool {
                const float n0 = fmaf(b.z, U2, fmaf(b.y, U1, fmaf(b.x, U0, U3)));
                const float n1 = fmaf(b.z, V2, fmaf(b.y, V1, fmaf(b.x, V0, V3)));
                const float iz = __frcp_approx(fmaf(b.z, C2, fmaf(b.y, C1, fmaf(b.x, C0, C3))));
                const float ux0 = n0 * iz, uy0 = n1 * iz;
                const float t = b.w * fabsf(iz);  // |iz|: with g = inf (AABB not safely in front) the radius must be +inf
                const float Rx = fmaf(t, fmaf(fabsf(ux0), Kg, KUg), m_hi), Ry = fmaf(t, fmaf(fabsf(uy0), Kg, KVg), m_hi);
                const float hi_x = fminf((ux0 + offx) + Rx, band_hi_x), lo_x = fmaxf((ux0 + offx_lo) - Rx, clamp_lo_x);
                const float hi_y = fminf((uy0 + offy) + Ry, band_hi_y), lo_y = fmaxf((uy0 + offy_lo) - Ry, clamp_lo_y);
                return (floorf(hi_x) < lo_x) || (floorf(hi_y) < lo_y);
            };
            // Levels 2 + 3 for one triangle given its projected vertices.  Level 2: can the approximate bbox hold a lattice
            // pixel of this band?  Level 3: if it holds exactly ONE, is that pixel clearly outside the triangle?  With
            // approximate vertices within d px of the exact ones, an edge function cross(Vj - Vi, P - Vi) differs from its
            // exact counterpart by less than 4.25 * d * D (D >= the largest distance involved; d = kEps3 scaled for near
            // vertices), so `sgn * E < -6 d D` together with `|2 * area| > 6 d D` (the orientation is then certain) implies a
            // strictly negative exact barycentric, i.e. the reference rejects the fragment (image_renderer.cuh:114-115).
            // Every rejecting comparison is false on NaN, so non-finite coordinates are kept (the exact path re-derives
            // everything); the test is symmetric in the vertex order.  Returns true = candidate for the exact path.
            auto tri_candidate = [&](float x0, float y0, float x1, float y1, float x2, float y2, float minx, float maxx,
                                     float miny, float maxy, float izmax) -> bool {
                const float lo_x = fmaxf(minx - m_lo3, clamp_lo_x), hi_x = fminf(maxx + m_hi, band_hi_x);
                const float lo_y = fmaxf(miny - m_lo3, clamp_lo_y), hi_y = fminf(maxy + m_hi, band_hi_y);
                const float px = ceilf(lo_x), py = ceilf(lo_y);  // first integer of [lo, hi] along each axis, if any
                if ((px > hi_x) || (py > hi_y)) return false;    // level 2: no lattice pixel
                const bool single = (px + 1.f > hi_x) && (py + 1.f > hi_y);
                const float ax = x1 - x0, ay = y1 - y0, bx = x2 - x0, by = y2 - y0;
                const float area2 = ax * by - bx * ay;
                const float sgn = area2 < 0.f ? -1.f : 1.f;
                const float e0 = ax * (py - y0) - ay * (px - x0);
                const float e1 = (x2 - x1) * (py - y1) - (y2 - y1) * (px - x1);
                const float e2 = -bx * (py - y2) + by * (px - x2);
                const float D = (maxx - minx) + (maxy - miny) + 1.2f * inv_s;
                const float tol = (6.f * kEps3 * inv_s) * fmaxf(1.f, kRefZ * izmax) * D;
                const float worst = fminf(fminf(e0 * sgn, e1 * sgn), e2 * sgn);
                return !(single && (fabsf(area2) > tol) && (worst < -tol) && (tol < 1e6f));
            };
            // One quad: four vertex transforms for two triangles (vertices 0,1 = shared edge, 2 = first triangle's own,
            // 3 = second's).  Bit 0 / 1 of the result = first / second triangle goes to the exact path.  A vertex near the
            // camera plane (the approximation's error bound grows like 1/z) sends both there untested.
            auto quad_candidates = [&](int q) -> unsigned {
                const float4* rec = reinterpret_cast<const float4*>(a.models.cl_recs + (int64_t)q * kQuadFloats);
                const float4 r0 = __ldg(rec), r1 = __ldg(rec + 1), r2 = __ldg(rec + 2);
                const float qx[4] = {r0.x, r0.w, r1.z, r2.y}, qy[4] = {r0.y, r1.x, r1.w, r2.z}, qz[4] = {r0.z, r1.y, r2.x, r2.w};
                float ux[4], uy[4], az[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const float x = qx[k], y = qy[k], z = qz[k];
                    const float n0 = fmaf(z, U2, fmaf(y, U1, fmaf(x, U0, U3)));
                    const float n1 = fmaf(z, V2, fmaf(y, V1, fmaf(x, V0, V3)));
                    const float iz = __frcp_approx(fmaf(z, C2, fmaf(y, C1, fmaf(x, C0, C3))));
                    ux[k] = fmaf(n0, iz, offx);
                    uy[k] = fmaf(n1, iz, offy);
                    az[k] = fabsf(iz);
                }
                // a NaN coordinate would be skipped by the min / max below (and its handling by the reference's select chain
                // is order dependent): both triangles go to the exact path, which re-derives everything
                const float sum = ((ux[0] + ux[1]) + (ux[2] + ux[3])) + ((uy[0] + uy[1]) + (uy[2] + uy[3]));
                const float iz01 = fmaxf(az[0], az[1]);
                if (!(sum == sum) || !(fmaxf(iz01, fmaxf(az[2], az[3])) < 1.0f / kNearZ)) return 3u;
                const float ex0 = fminf(ux[0], ux[1]), ex1 = fmaxf(ux[0], ux[1]);
                const float ey0 = fminf(uy[0], uy[1]), ey1 = fmaxf(uy[0], uy[1]);
                unsigned r = 0u;
                if (tri_candidate(ux[0], uy[0], ux[1], uy[1], ux[2], uy[2], fminf(ex0, ux[2]), fmaxf(ex1, ux[2]),
                                  fminf(ey0, uy[2]), fmaxf(ey1, uy[2]), fmaxf(iz01, az[2])))
                    r |= 1u;
                if (tri_candidate(ux[0], uy[0], ux[1], uy[1], ux[3], uy[3], fminf(ex0, ux[3]), fmaxf(ex1, ux[3]),
                                  fminf(ey0, uy[3]), fmaxf(ey1, uy[3]), fmaxf(iz01, az[3])))
                    r |= 2u;
                return r;
            };
            const unsigned lt_mask = (1u << lane) - 1u;
            // levels 2 + 3 of `count` <= 32 pending quads, one lane each; the survivors go to the CTA's queue with one
            // atomic per call.  A slot that merely repeats its quad's other triangle (kPermInvalid) is never a candidate.
            auto test_pending = [&](int first, int count) {
                unsigned hits = 0;
                int ea = 0, eb = 0;
                if (lane < count) {
                    const int q = IDX_OK(first + lane, kPendCap) ? wpend[first + lane] : 2 * cl_low;
                    const uint32_t perm = __ldg(a.models.cl_perm + (q >> 1)) >> (12 * (q & 1));
                    const uint32_t pa = perm & 63u, pb = (perm >> 6) & 63u;
                    hits = quad_candidates(q);
                    if (pa == kPermInvalid) hits &= ~1u;
                    if (pb == kPermInvalid) hits &= ~2u;
                    ea = (2 * q) | (int)(pa << kSlotBits);
                    eb = (2 * q + 1) | (int)(pb << kSlotBits);
                }
                const unsigned b0 = __ballot_sync(0xffffffffu, hits & 1u), b1 = __ballot_sync(0xffffffffu, hits & 2u);
                const int n0 = __popc(b0), total = n0 + __popc(b1);
                if (lane == 0) { STAT(1, count); STAT(2, count); STAT(3, total); STAT(9, 1); }
                int slot = 0;
                if (lane == 0 && total) slot = smem_add_elected(&s_qcount, total);
                slot = __shfl_sync(0xffffffffu, slot, 0);
                if ((hits & 1u) && IDX_OK(slot + __popc(b0 & lt_mask), kQueueCap)) queue[slot + __popc(b0 & lt_mask)] = ea;
                if ((hits & 2u) && IDX_OK(slot + n0 + __popc(b1 & lt_mask), kQueueCap)) queue[slot + n0 + __popc(b1 & lt_mask)] = eb;
            };

            // Warps claim groups of kClaim quads dynamically (no block barrier inside pass A); a warp stops claiming when
            // the candidate queue could overflow.
            const int q_low = 2 * cl_low, q_high = 2 * cl_high;
            const int n_groups = (q_high - q_low + kClaim - 1) / kClaim;
            if (tid == 0) s_next = 0;
            __syncthreads();
            for (;;) {
                int npend = 0;  // warp-uniform
                for (;;) {
                    int g = 0x7fffffff;
                    if (lane == 0 && *(volatile int*)&s_qcount <= kQueueCap - kWarps * kClaimMaxTris) g = smem_add_elected(&s_next, 1);
                    g = __shfl_sync(0xffffffffu, g, 0);
                    if (g >= n_groups) break;
                    const int qb = q_low + g * kClaim + lane;
                    float4 bnd[kLaneQuads];
#pragma unroll
                    for (int k = 0; k < kLaneQuads; ++k) bnd[k] = __ldg(a.models.q_bounds + min(qb + 32 * k, q_high - 1));
                    // survivors of the claim's kClaim spheres go to the warp's list with ONE compaction (a lane's kept
                    // quads, then a shuffle scan over the lanes) instead of a ballot + popc + store per 32 spheres:
                    // C2 1.239 -> 1.230 ms, C3 1.292 -> 1.276 ms.  The list order changes, the set does not.
                    unsigned keepmask = 0u;
#pragma unroll
                    for (int k = 0; k < kLaneQuads; ++k) {
                        const int q = qb + 32 * k;
                        if ((q < q_high) && !sphere_rejects(bnd[k])) keepmask |= 1u << k;
                    }
                    {
                        const int cnt = __popc(keepmask);
                        int incl = cnt;
#pragma unroll
                        for (int o = 1; o < 32; o <<= 1) {
                            const int v = __shfl_up_sync(0xffffffffu, incl, o);
                            if (lane >= o) incl += v;
                        }
                        int slot = npend + incl - cnt;
#pragma unroll
                        for (int k = 0; k < kLaneQuads; ++k)
                            if ((keepmask >> k) & 1u) { if (IDX_OK(slot, kPendCap)) wpend[slot] = qb + 32 * k; ++slot; }
                        npend += __shfl_sync(0xffffffffu, incl, 31);
                    }
                    if (lane == 0) { STAT(0, min(kClaim, q_high - (q_low + g * kClaim))); }
                    __syncwarp();
                    while (npend >= 32) {
                        npend -= 32;
                        test_pending(npend, 32);
                    }
                    __syncwarp();
                }
                if (npend > 0) test_pending(0, npend);
                __syncthreads();
                const int qn = s_qcount;
                const bool all_claimed = *(volatile int*)&s_next >= n_groups;
                // pass B: exact rasterisation of the queued triangles, one per thread (round 0); triangles covering many
                // lattice pixels are deferred to the cooperative round 1, every thread on the same triangle.
#pragma unroll 1
                for (int round = 0; round < 2; ++round) {
                    const bool coop = round == 1;
                    const int n_items = coop ? min(s_bigcount, kBigCap) : qn;
                    const int* list = coop ? s_big : queue;
                    const int e_step = coop ? 1 : kThreads;
                    for (int e = coop ? 0 : tid; e < n_items; e += e_step) raster_tri(list[e], coop);
                    __syncthreads();
                }
                if (tid == 0) { s_qcount = 0; s_bigcount = 0; }
                __syncthreads();
                if (tid == 0) STAT(10, 1);
                if (all_claimed) break;
            }
            if (tid == 0) { STAT(11, 1); STAT(12, band_cells); }

            if (MODE == MODE_IMAGE) {
                // ---- 3'. resolve the band into the output images (or into the shared key image) -------------
                for (int c = tid; c < band_cells; c += kThreads) {
                    zkey k = zbuf[c];
                    int lx = rx0 + c % tile_w, ly = by0 + c / tile_w;
                    if (a.single_keys) {
                        if (k != Z::empty()) atomicMin(a.single_keys + (size_t)ly * Lw + lx, (unsigned long long)k);
                        continue;
                    }
                    int32_t d = 0;
                    uint32_t rgb = 0;
                    if (k != Z::empty()) {
                        d = Z::depth(k);
                        int u = lx * s, v = ly * s;
                        int32_t src = __ldg(a.scene.depth + v * W + u);
                        int src_label = (use_label && a.scene.label) ? (int)__ldg(a.scene.label + v * W + u) : 0;
                        if (a.tree_total) {
                            // image_renderer.cuh:163-206 on the pixel's minimum depth: the gate of the occlusion rule,
                            // then "source hides render" (cleared, and counted as clutter: the reference's extra
                            // `source <= new_depth - 5` reads the pixel AFTER it was set to INT_MAX, so it always
                            // holds) or "render hides source"
                            ++tree_total_local;
                            const int ad = abs(d - src);
                            const bool gate = use_label ? (pose_label != src_label - 1 && ad >= 1)
                                                        : ((float)ad > a.occlusion_threshold);
                            if (gate && src > 0) {
                                if (d > src) {
                                    tree_flags_local |= 2;
                                    ++tree_clutter_local;
                                } else {
                                    tree_flags_local |= 1;
                                }
                            }
                        }
                        if (source_occludes(d, src, use_label, pose_label, src_label, a.occlusion_threshold)) d = 0;
                        else rgb = __ldg(a.models.tri_color + Z::tri(k));
                        if (d == INT_MAX) d = 0;  // max2zero (image_renderer.cuh:324-333,465-466)
                    }
                    if (!a.img_depth) continue;
                    size_t o = ((size_t)pose_i * Lh + ly) * Lw + lx;
                    a.img_depth[o] = d;
                    if (a.img_r) { a.img_r[o] = (uint8_t)rgb; a.img_g[o] = (uint8_t)(rgb >> 8); a.img_b[o] = (uint8_t)(rgb >> 16); }
                }
            } else {
#ifdef PERCH_COST_SMEM_OBS
                // A/B variant: the pose's label range of the sorted observed cloud -> shared memory with one bulk copy
                // (cp.async.bulk + mbarrier), then a linear scan from shared memory instead of the grid query
                float4* s_obs = reinterpret_cast<float4*>((reinterpret_cast<uintptr_t>(mask + mask_words) + 15) & ~(uintptr_t)15);
                __shared__ int s_obs_n, s_obs_lb;
                __shared__ __align__(8) unsigned long long s_bar;
                if (tid == 0) {
                    int lb = 0, le = a.grid.n_obs;
                    if (use_label) {
                        const bool known = pose_label >= 0 && pose_label < a.n_label_ranges;
                        lb = known ? a.label_range[2 * pose_label] : 0;
                        le = known ? a.label_range[2 * pose_label + 1] : 0;
                    }
                    s_obs_lb = lb;
                    s_obs_n = (le - lb) <= PERCH_COST_SMEM_OBS ? le - lb : -1;  // too large: the grid query
                    const unsigned bar = (unsigned)__cvta_generic_to_shared(&s_bar);
                    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
                    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
                    if (s_obs_n > 0) {
                        const unsigned bytes = (unsigned)s_obs_n * 16u;
                        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
                        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                                     ::"r"((unsigned)__cvta_generic_to_shared(s_obs)), "l"(a.obs_sorted + lb), "r"(bytes), "r"(bar) : "memory");
                    }
                }
                __syncthreads();
                if (s_obs_n > 0) {
                    const unsigned bar = (unsigned)__cvta_generic_to_shared(&s_bar);
                    unsigned done = 0;
                    while (!done)
                        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }"
                                     : "=r"(done) : "r"(bar) : "memory");
                }
#endif
                // ---- 3. cost over the band's lattice pixels ---------------------------------------------------
                // 3a: covered pixels that survive the source-occlusion rule are compacted into the (now idle) candidate
                // queue, so that the nearest-neighbour queries below run with every lane busy
                const bool compact = band_cells <= kQueueCap;
                if (compact) {
                    if (tid == 0) s_qcount = 0;
                    __syncthreads();
                    for (int c0 = 0; c0 < band_cells; c0 += kThreads) {
                        const int c = c0 + tid;
                        bool live = false;
                        if (c < band_cells) {
                            const zkey k = zbuf[c];
                            if (k != Z::empty()) {
                                const int32_t d = Z::depth(k);
                                const int u = (rx0 + c % tile_w) * s, v = (by0 + c / tile_w) * s;
                                const int32_t src = __ldg(a.scene.depth + v * W + u);
                                const int src_label = (use_label && a.scene.label) ? (int)__ldg(a.scene.label + v * W + u) : 0;
                                // compute_point_clouds.cuh:64,123: only positive depths become points
                                live = !source_occludes(d, src, use_label, pose_label, src_label, a.occlusion_threshold) && d > 0;
                            }
                        }
                        const unsigned lb = __ballot_sync(0xffffffffu, live);
                        int base = 0;
                        if (lane == 0 && lb) base = smem_add_elected(&s_qcount, __popc(lb));
                        base = __shfl_sync(0xffffffffu, base, 0);
                        if (live && IDX_OK(base + __popc(lb & ((1u << lane) - 1u)), kQueueCap)) queue[base + __popc(lb & ((1u << lane) - 1u))] = c;
                    }
                    __syncthreads();
                }
                const int n_live = compact ? s_qcount : band_cells;
                for (int e = tid; e < n_live; e += kThreads) {
                    const int c = compact ? queue[e] : e;
                    zkey k = zbuf[c];
                    int32_t d = Z::depth(k);
                    int lx = rx0 + c % tile_w, ly = by0 + c / tile_w;
                    int u = lx * s, v = ly * s;
                    if (!compact) {
                        if (k == Z::empty()) continue;
                        int32_t src = __ldg(a.scene.depth + v * W + u);
                        int src_label = (use_label && a.scene.label) ? (int)__ldg(a.scene.label + v * W + u) : 0;
                        if (source_occludes(d, src, use_label, pose_label, src_label, a.occlusion_threshold)) continue;
                        if (d <= 0) continue;  // compute_point_clouds.cuh:64,123
                    }
                    ++n_local;
                    float qx, qy, qz;
                    back_project(u, v, d, a.cam.cx, a.cam.cy, a.cam.fx, a.cam.fy, a.cam.depth_factor, qx, qy, qz);
                    float best; int ref;
#ifdef PERCH_COST_SMEM_OBS
                    if (s_obs_n >= 0) {
                        best = FLT_MAX; ref = -1;
                        for (int j = 0; j < s_obs_n; ++j) {
                            const float4 p = s_obs[j];
                            const float dd = sqdist3(qx, qy, qz, p.x, p.y, p.z);
                            if (dd < best) { best = dd; ref = s_obs_lb + j; }
                        }
                    } else
#endif
                    nn_query(a.grid, qx, qy, qz, use_label, pose_label, best, ref);
                    if (best > a.res2) { ++c_local; continue; }  // compute_costs.cuh:206-209
                    if (COLOR) {
                        uint32_t t = Z::tri(k);
                        const float* lr = a.tri_lab + 3 * (size_t)t;
                        const float* lo = a.obs_lab + 3 * (size_t)ref;
                        double cd = ciede2000_ref(lo[0], lo[1], lo[2], lr[0], lr[1], lr[2]);
                        if (cd > a.color_thr) { ++c_local; continue; }  // compute_costs.cuh:231-234
                    }
                    if (IDX_OK(ref >> 5, mask_words)) atomicOr(&mask[ref >> 5], 1u << (ref & 31));  // explained[pose][nn] = 1 (compute_costs.cuh:238,244,268)
                }
            }
            __syncthreads();
        }
    }

    if (MODE == MODE_IMAGE) {
        // every output pixel outside the pose tile is empty: the kernel writes the whole image, no memset pass
        if (!a.single_keys && a.img_depth) {
            const size_t img0 = (size_t)pose_i * Lh * Lw;
            const int n_px = Lw * Lh;
            // groups of four pixels that lie entirely outside the tile are cleared with one 16-byte store (depth) and
            // three 4-byte stores (colour planes); requires rows and image bases that keep the groups aligned
            const bool wide = (Lw % 4 == 0) && (img0 % 4 == 0) && ((reinterpret_cast<uintptr_t>(a.img_depth) & 15) == 0) &&
                              (!a.img_r || (((reinterpret_cast<uintptr_t>(a.img_r) | reinterpret_cast<uintptr_t>(a.img_g) |
                                              reinterpret_cast<uintptr_t>(a.img_b)) & 3) == 0));
            if (wide) {
                const int Lw4 = Lw >> 2;
                for (int g = tid; g < (n_px >> 2); g += kThreads) {
                    const int ly = g / Lw4, lx = (g - ly * Lw4) << 2;
                    const size_t o = img0 + (size_t)g * 4;
                    const bool row_in = !tile_empty && ly >= ry0 && ly <= ry1;
                    if (!row_in || lx + 3 < rx0 || lx > rx1) {
                        *reinterpret_cast<int4*>(a.img_depth + o) = make_int4(0, 0, 0, 0);
                        if (a.img_r) {
                            *reinterpret_cast<uint32_t*>(a.img_r + o) = 0u;
                            *reinterpret_cast<uint32_t*>(a.img_g + o) = 0u;
                            *reinterpret_cast<uint32_t*>(a.img_b + o) = 0u;
                        }
                    } else {
                        for (int k = 0; k < 4; ++k) {
                            if (lx + k >= rx0 && lx + k <= rx1) continue;
                            a.img_depth[o + k] = 0;
                            if (a.img_r) { a.img_r[o + k] = 0; a.img_g[o + k] = 0; a.img_b[o + k] = 0; }
                        }
                    }
                }
            } else {
                for (int i = tid; i < n_px; i += kThreads) {
                    int lx = i % Lw, ly = i / Lw;
                    if (!tile_empty && lx >= rx0 && lx <= rx1 && ly >= ry0 && ly <= ry1) continue;
                    size_t o = img0 + i;
                    a.img_depth[o] = 0;
                    if (a.img_r) { a.img_r[o] = 0; a.img_g[o] = 0; a.img_b[o] = 0; }
                }
            }
        }
        if (a.tree_total) {
#pragma unroll
            for (int o = 16; o >= 1; o >>= 1) {
                tree_flags_local |= __shfl_xor_sync(0xffffffffu, tree_flags_local, o);
                tree_clutter_local += __shfl_xor_sync(0xffffffffu, tree_clutter_local, o);
                tree_total_local += __shfl_xor_sync(0xffffffffu, tree_total_local, o);
            }
            if ((tid & 31) == 0) {
                if (tree_flags_local) atomicOr(&s_n, tree_flags_local);
                if (tree_clutter_local) atomicAdd(&s_c, tree_clutter_local);
                if (tree_total_local) atomicAdd(&s_e, tree_total_local);
            }
            __syncthreads();
            if (tid == 0) {
                const float clutter = (float)s_c, total = (float)s_e;
                if (a.tree_occluded) a.tree_occluded[pose_i] = s_n & 1;
                if (a.tree_occluded_other) a.tree_occluded_other[pose_i] = (s_n >> 1) & 1;
                // thrust::divides then multiplies by 100 (image_renderer.cuh:467-482); 0/0 = NaN like the reference
                if (a.tree_clutter) a.tree_clutter[pose_i] = __fmul_rn(__fdiv_rn(clutter, total), 100.f);
                a.tree_total[pose_i] = total;
            }
        }
        return;
    }

    // ---- reduce counters ---------------------------------------------------------------------------------
    int e_local = 0;
    for (int w = tid; w < mask_words; w += kThreads) e_local += __popc(mask[w]);
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) {
        n_local += __shfl_xor_sync(0xffffffffu, n_local, o);
        c_local += __shfl_xor_sync(0xffffffffu, c_local, o);
        e_local += __shfl_xor_sync(0xffffffffu, e_local, o);
    }
    if ((tid & 31) == 0) {
        if (n_local) atomicAdd(&s_n, n_local);
        if (c_local) atomicAdd(&s_c, c_local);
        if (e_local) atomicAdd(&s_e, e_local);
    }
    __syncthreads();
    if (tid == 0) {
        const int P = a.n_poses;
        float n = (float)s_n, c = (float)s_c, e = (float)s_e;
        // compute_costs.cuh:364-378: explained = n - c; cost = (n == 0) ? -1 : c / n; cost = (cost == -1) ? -1 : cost*100
        float rendered_explained = __fsub_rn(n, c);
        float rc = (n == 0.f) ? -1.f : __fdiv_rn(c, n);
        rc = (rc == -1.f) ? -1.f : __fmul_rn(rc, 100.f);
        a.out_costs[pose_i] = rc;
        if (a.calc_observed) {
            // compute_costs.cuh:407-446: diff = explained - explained_obs; observed = ((tot - explained_obs)/tot)*100
            float tot = a.pose_obs_total ? a.pose_obs_total[pose_i] : 0.f;
            a.out_costs[P + pose_i] = __fmul_rn(__fdiv_rn(__fsub_rn(tot, e), tot), 100.f);
            a.out_costs[2 * P + pose_i] = __fsub_rn(rendered_explained, e);
        }
        if (a.out_counts) {
            a.out_counts[pose_i] = s_n;
            a.out_counts[P + pose_i] = s_c;
            a.out_counts[2 * P + pose_i] = s_e;
        }
    }
}

// keys -> images for single_result_image mode
__global__ void resolve_single_kernel(PoseArgs a) {
    const int W = a.cam.width, H = a.cam.height;
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= W * H) return;
    unsigned long long k = a.single_keys[i];
    int32_t d = 0;
    uint32_t rgb = 0;
    if (k != ~0ull) {
        d = ZCell<true>::depth(k);
        const bool use_label = a.pose_label != nullptr;
        int pose_label = use_label ? a.pose_label[0] : 0;
        int src_label = (use_label && a.scene.label) ? (int)a.scene.label[i] : 0;
        if (source_occludes(d, a.scene.depth[i], use_label, pose_label, src_label, a.occlusion_threshold)) d = 0;
        else rgb = a.models.tri_color[ZCell<true>::tri(k)];
        if (d == INT_MAX) d = 0;
    }
    a.img_depth[i] = d;
    if (a.img_r) { a.img_r[i] = (uint8_t)rgb; a.img_g[i] = (uint8_t)(rgb >> 8); a.img_b[i] = (uint8_t)(rgb >> 16); }
}

__global__ void fill_keys_kernel(unsigned long long* keys, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) keys[i] = ~0ull;
}

}  // namespace

size_t pose_kernel_smem_bytes(const PoseArgs& a, bool image_mode, bool color) {
    size_t b = (size_t)a.tile_cap * ((color || image_mode) ? 8 : 4) + (image_mode ? 0 : (size_t)a.mask_words * 4) +
               (size_t)kQueueCap * 4;
#ifdef PERCH_COST_SMEM_OBS
    if (!image_mode) b += (size_t)PERCH_COST_SMEM_OBS * sizeof(float4) + 16;
#endif
    return b + (size_t)kWarps * kPendCap * sizeof(int);
}

template <int MODE, bool COLOR>
static cudaError_t launch_pose_kernel(const PoseArgs& a, cudaStream_t stream) {
    if (a.n_poses <= 0) return cudaSuccess;
    size_t smem = pose_kernel_smem_bytes(a, MODE == MODE_IMAGE, COLOR);
    // The attribute is per function and per device, not per context: always set it to the device's opt-in limit (the same
    // value from every thread and context) instead of this launch's size, so that concurrent contexts with different
    // sizes cannot lower it under each other's launches.  The caller has checked smem against the limit.
    int dev = 0, limit = 0;
    cudaError_t err = cudaGetDevice(&dev);
    if (err == cudaSuccess) err = cudaDeviceGetAttribute(&limit, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    if (err != cudaSuccess) return err;
    cudaFuncAttributes fa;
    if ((err = cudaFuncGetAttributes(&fa, pose_kernel<MODE, COLOR>)) != cudaSuccess) return err;
    limit -= (int)fa.sharedSizeBytes;  // the opt-in limit covers static + dynamic shared memory
    if (smem > (size_t)limit) return cudaErrorInvalidValue;
    static int s_set[64][2][2];  // [device][MODE][COLOR]: attribute already raised (benign race: same value)
    if (dev < 0 || dev >= 64 || !s_set[dev][MODE][COLOR]) {
        err = cudaFuncSetAttribute(pose_kernel<MODE, COLOR>, cudaFuncAttributeMaxDynamicSharedMemorySize, limit);
        if (err != cudaSuccess) return err;
        if (dev >= 0 && dev < 64) s_set[dev][MODE][COLOR] = 1;
    }
    pose_kernel<MODE, COLOR><<<a.n_poses, kThreads, smem, stream>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_fused_cost(const PoseArgs& a, bool color, cudaStream_t stream) {
    return color ? launch_pose_kernel<MODE_COST, true>(a, stream) : launch_pose_kernel<MODE_COST, false>(a, stream);
}

cudaError_t launch_raster_images(const PoseArgs& a, cudaStream_t stream) {
    return launch_pose_kernel<MODE_IMAGE, true>(a, stream);
}

cudaError_t launch_single_clear(unsigned long long* keys, int n, cudaStream_t stream) {
    fill_keys_kernel<<<(n + 255) / 256, 256, 0, stream>>>(keys, n);
    return cudaGetLastError();
}

cudaError_t launch_resolve_single(const PoseArgs& a, cudaStream_t stream) {
    int n = a.cam.width * a.cam.height;
    resolve_single_kernel<<<(n + 255) / 256, 256, 0, stream>>>(a);
    return cudaGetLastError();
}

}  // namespace perch

#ifdef PERCH_STATS
extern "C" int perch_debug_stats(unsigned long long* out16, int reset) {
    if (out16 && cudaMemcpyFromSymbol(out16, perch::g_stats, sizeof(perch::g_stats)) != cudaSuccess) return 1;
    if (reset) {
        unsigned long long z[16] = {0};
        if (cudaMemcpyToSymbol(perch::g_stats, z, sizeof(z)) != cudaSuccess) return 1;
    }
    return 0;
}
#endif
