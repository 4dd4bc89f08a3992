// Full-resolution rasteriser (stages RENDER / DEBUG, scene composition, perch_tree_flags): image_render
// (image_renderer.cuh:336-496) for EVERY pixel.
//
// The per-pose kernel of pose_kernels.cu is built around the stride lattice: with stride 1 nothing can be culled, every
// triangle reaches its exact path, and the 64-bit shared z-buffer leaves 3 CTAs per SM
// (profiles/r1_render_kernel_v1.md: 8 % of the HBM write roofline).  Here the work is spread over (pose, 128 quads) CTAs
// instead of one CTA per pose:
//   raster_quads_kernel  one thread per quad: FOUR exact vertex transforms for two triangles, the reference's bbox and
//                        inside test (exact_math.cuh); pixels that pass are queued per warp and their perspective
//                        depths (three IEEE divisions) computed with every lane busy; triangles of more than a few
//                        pixels are walked by the whole warp; z-test with a 64-bit atomicMin on an order-preserving
//                        (depth, triangle) key image in global memory (L2 atomics);
//   raster_big_kernel    triangles covering thousands of pixels, deferred by the first kernel, one warp each;
//   resolve_full_kernel  keys -> int32 depth + RGB with the source-occlusion rule, four pixels per thread (16-byte
//                        stores), and optionally the per-pose vectors of the USE_TREE / USE_CLUTTER branches.
// Canonical semantics as everywhere: minimum depth over fragments, lowest original triangle index
// at that depth wins the colour, occlusion rule on the final depth.
#include <cuda_runtime.h>

#include <algorithm>
#include <climits>
#include <cstdio>
#include <cstdlib>

#include "clusters.hpp"
#include "context.hpp"
#include "kernels.cuh"

namespace perch {

namespace {

constexpr int kFullThreads = 128;
constexpr int kBigThreads = 256;

__device__ __forceinline__ unsigned long long make_key(int32_t d, uint32_t id) {
    return ((unsigned long long)((uint32_t)d ^ 0x80000000u) << 32) | id;
}

struct PixelRect {
    int x0, x1, r0, r1;  // inclusive pixel columns and RASTER rows (image row = H-1-r); empty when x0 > x1 or r0 > r1
};

// The pixels the reference's loops visit (image_renderer.cuh:110-111): P from size_t(bboxmin + 0.5f) while P <= bboxmax.
__device__ __forceinline__ PixelRect pixel_rect_of(float bminx, float bmaxx, float bminy, float bmaxy, int W, int H) {
    PixelRect r{1, 0, 1, 0};
    if (!(bmaxx >= 0.f) || !(bmaxy >= 0.f)) return r;  // also rejects NaN upper bounds (loop condition false)
    const unsigned long long p0 = (unsigned long long)__fadd_rn(bminx, 0.5f);  // NaN -> 0
    const unsigned long long p1 = (unsigned long long)__fadd_rn(bminy, 0.5f);
    if (p0 > (unsigned long long)(W - 1) || p1 > (unsigned long long)(H - 1)) return r;
    r.x0 = (int)p0; r.x1 = (int)bmaxx;  // <= W-1 / H-1 by the clamp of bbox_of
    r.r0 = (int)p1; r.r1 = (int)bmaxy;
    return r;
}

struct Tri3 {
    ScreenVertex a, b, c;
};

__device__ __forceinline__ ScreenVertex pick(const ScreenVertex v[4], uint32_t i) {
    ScreenVertex o;
    o.sx = i == 0 ? v[0].sx : i == 1 ? v[1].sx : i == 2 ? v[2].sx : v[3].sx;
    o.sy = i == 0 ? v[0].sy : i == 1 ? v[1].sy : i == 2 ? v[2].sy : v[3].sy;
    o.z = i == 0 ? v[0].z : i == 1 ? v[1].z : i == 2 ? v[2].z : v[3].z;
    return o;
}

// ---- fragments in two steps ------------------------------------------------------------------------------------------
// The inside test (image_renderer.cuh:44-57,114-118) is cheap and fails for most bbox pixels of small triangles; the
// perspective depth (three IEEE divisions, 119-129) is expensive.  Lanes run the test on their own triangles and queue
// the few pixels that pass; the warp then computes the depths with every lane busy.  Same operations on the same values
// as fragment_depth (exact_math.cuh), only regrouped.
constexpr int kFragCap = 128;  // queue entries per warp; drained when fewer than 32 are free

struct FragQueue {  // SoA, one per warp
    float beta[kFragCap], gamma[kFragCap], z0[kFragCap], z1[kFragCap], z2[kFragCap];
    uint32_t pixel[kFragCap], id[kFragCap];
};

__device__ __forceinline__ bool inside_test(const TriSetup& t, float P0f, float P1f, float& beta, float& gamma) {
    const float dy = __fsub_rn(P1f, t.A1);
    const float r0 = __fmul_rn(t.e_c0, dy);
    const float r1 = __fmul_rn(dy, t.e_b0);
    const float dx = __fsub_rn(P0f, t.A0);
    beta = __fmul_rn(t.base_inv, __fmul_rn(__fmaf_rn(t.e_c1, -dx, r0), 0.5f));
    gamma = __fmul_rn(t.base_inv, __fmul_rn(__fmaf_rn(t.e_b1, dx, -r1), 0.5f));
    const float alpha = __fsub_rn(__fsub_rn(1.0f, beta), gamma);
    const float mn = fminf(gamma, fminf(beta, alpha));
    const float mx = fmaxf(fmaxf(alpha, beta), gamma);
    return !(mn < 0.0f || mx > 1.0f);
}

__device__ __forceinline__ void drain(FragQueue& q, int count, int lane, unsigned long long* keys) {
    __syncwarp();
    for (int e = lane; e < count; e += 32) {
        const float beta = q.beta[e], gamma = q.gamma[e];
        const float alpha = __fsub_rn(__fsub_rn(1.0f, beta), gamma);
        const float num = __fadd_rn(gamma, __fadd_rn(beta, alpha));
        const float den = __fadd_rn(__fdiv_rn(gamma, q.z2[e]), __fadd_rn(__fdiv_rn(beta, q.z1[e]), __fdiv_rn(alpha, q.z0[e])));
        const int32_t d = (int32_t)__fadd_rn(__fdiv_rn(num, den), 0.5f);
        atomicMin(keys + q.pixel[e], make_key(d, q.id[e]));
    }
    __syncwarp();
}

// Warp-uniform call: lanes with `inside` append their fragment; returns the new (uniform) count.
__device__ __forceinline__ int append(FragQueue& q, int count, int lane, bool inside, float beta, float gamma,
                                      const TriSetup& t, uint32_t pixel, uint32_t id, unsigned long long* keys) {
    const unsigned m = __ballot_sync(0xffffffffu, inside);
    if (m == 0) return count;
    if (inside) {
        const int s = count + __popc(m & ((1u << lane) - 1u));
        q.beta[s] = beta; q.gamma[s] = gamma; q.z0[s] = t.z0; q.z1[s] = t.z1; q.z2[s] = t.z2;
        q.pixel[s] = pixel; q.id[s] = id;
    }
    count += __popc(m);
    if (count > kFragCap - 32) {
        drain(q, count, lane, keys);
        count = 0;
    }
    return count;
}

__global__ void __launch_bounds__(kFullThreads) raster_quads_kernel(FullRasterArgs a) {
    __shared__ Mat4 s_pose;
    __shared__ FragQueue s_queue[kFullThreads / 32];
    const int pose_i = blockIdx.y, tid = threadIdx.x, lane = tid & 31;
    const int model = a.pose_model_map[pose_i];
    if (model < 0 || model >= a.models.n_models) return;
    const int q_low = 2 * a.models.cl_low[model], q_high = 2 * a.models.cl_high[model];
    if (q_low + (int)blockIdx.x * kFullThreads >= q_high) return;  // whole CTA past this model's quads
    if (tid < 16) reinterpret_cast<float*>(&s_pose)[tid] = reinterpret_cast<const float*>(a.poses + pose_i)[tid];
    __syncthreads();
    const int qi = q_low + blockIdx.x * kFullThreads + tid;
    const int W = a.cam.width, H = a.cam.height;
    unsigned long long* keys = a.keys + (a.single ? 0 : (size_t)pose_i * W * H);
    FragQueue& q = s_queue[tid >> 5];
    int cells[2] = {0, 0};  // bbox pixels of the lane's two triangles (0: nothing to draw, or deferred)
    PixelRect rect[2] = {{1, 0, 1, 0}, {1, 0, 1, 0}};
    TriSetup ts[2];
    uint32_t id[2] = {0, 0};
    if (qi < q_high) {
        const uint32_t perm2 = (__ldg(a.models.cl_perm + (qi >> 1)) >> (12 * (qi & 1))) & 0xFFFu;
        const float wf = (float)W, hf = (float)H, cmaxx = (float)(W - 1), cmaxy = (float)(H - 1);
        const float4* rec = reinterpret_cast<const float4*>(a.models.cl_recs + (int64_t)qi * kQuadFloats);
        const float4 r0 = __ldg(rec), r1 = __ldg(rec + 1), r2 = __ldg(rec + 2);
        ScreenVertex v[4];
        v[0] = transform_vertex(s_pose, a.cam.proj, wf, hf, r0.x, r0.y, r0.z);
        v[1] = transform_vertex(s_pose, a.cam.proj, wf, hf, r0.w, r1.x, r1.y);
        v[2] = transform_vertex(s_pose, a.cam.proj, wf, hf, r1.z, r1.w, r2.x);
        v[3] = transform_vertex(s_pose, a.cam.proj, wf, hf, r2.y, r2.z, r2.w);
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const uint32_t perm = (perm2 >> (6 * j)) & 63u;
            if (perm == kPermInvalid) continue;
            const ScreenVertex A = pick(v, perm & 3u), B = pick(v, (perm >> 2) & 3u), C = pick(v, perm >> 4);
            float bminx, bmaxx, bminy, bmaxy;
            bbox_of(A, B, C, cmaxx, cmaxy, bminx, bmaxx, bminy, bmaxy);
            rect[j] = pixel_rect_of(bminx, bmaxx, bminy, bmaxy, W, H);
            if (rect[j].x0 > rect[j].x1 || rect[j].r0 > rect[j].r1) continue;
            const int t = 2 * qi + j;  // triangle slot
            const long long n = (long long)(rect[j].x1 - rect[j].x0 + 1) * (rect[j].r1 - rect[j].r0 + 1);
            if (n > a.big_cells) {
                const int slot = atomicAdd(a.big_count, 1);
                if (slot < a.big_cap) {
                    a.big_list[slot] = make_int2(pose_i, t | (int)(perm << 26));
                    continue;
                }
                // list full: the warp rasterises it below after all (slow for a huge triangle, correct)
            }
            setup_edges(A, B, C, ts[j]);
            id[j] = a.key_pose_id ? (uint32_t)pose_i : (uint32_t)__ldg(a.models.cl_orig + t);
            cells[j] = (int)(n > 0x7fffffffLL ? 0x7fffffffLL : n);  // W*H <= 2^26 (perch_set_camera)
        }
    }
    if (a.rect_min) {
        // rectangle of key pixels this warp's triangles can touch (deferred ones included), image rows: H-1-r
        int mnx = INT_MAX, mny = INT_MAX, mxx = INT_MIN, mxy = INT_MIN;
#pragma unroll
        for (int j = 0; j < 2; ++j)
            if (qi < q_high && rect[j].x0 <= rect[j].x1 && rect[j].r0 <= rect[j].r1) {
                mnx = min(mnx, rect[j].x0); mxx = max(mxx, rect[j].x1);
                mny = min(mny, H - 1 - rect[j].r1); mxy = max(mxy, H - 1 - rect[j].r0);
            }
        mnx = __reduce_min_sync(0xffffffffu, mnx); mny = __reduce_min_sync(0xffffffffu, mny);
        mxx = __reduce_max_sync(0xffffffffu, mxx); mxy = __reduce_max_sync(0xffffffffu, mxy);
        if (lane == 0 && mnx <= mxx) {
            const int r = a.single ? 0 : pose_i;
            atomicMin(a.rect_min + 2 * r, mnx); atomicMin(a.rect_min + 2 * r + 1, mny);
            atomicMax(a.rect_max + 2 * r, mxx); atomicMax(a.rect_max + 2 * r + 1, mxy);
        }
    }
    int count = 0;
#pragma unroll
    for (int j = 0; j < 2; ++j) {
        // (1) triangles of a few pixels: every lane walks its own bbox
        const int mine = cells[j] <= a.lane_cells ? cells[j] : 0;
        const int trips = __reduce_max_sync(0xffffffffu, mine);
        int p0 = rect[j].x0, p1 = rect[j].r0;
        for (int it = 0; it < trips; ++it) {
            float beta = 0.f, gamma = 0.f;
            bool in = false;
            uint32_t pixel = 0;
            if (it < mine) {
                in = inside_test(ts[j], (float)p0, (float)p1, beta, gamma);
                pixel = (uint32_t)((H - 1 - p1) * W + p0);  // image_renderer.cuh:127-128
                if (++p0 > rect[j].x1) { p0 = rect[j].x0; ++p1; }
            }
            count = append(q, count, lane, in, beta, gamma, ts[j], pixel, id[j], keys);
        }
        // (2) larger ones: the warp takes them one at a time, 32 pixels per step
        unsigned todo = __ballot_sync(0xffffffffu, cells[j] > a.lane_cells);
        while (todo) {
            const int src = __ffs(todo) - 1;
            todo &= todo - 1;
            TriSetup t;
            t.A0 = __shfl_sync(0xffffffffu, ts[j].A0, src); t.A1 = __shfl_sync(0xffffffffu, ts[j].A1, src);
            t.e_c0 = __shfl_sync(0xffffffffu, ts[j].e_c0, src); t.e_b1 = __shfl_sync(0xffffffffu, ts[j].e_b1, src);
            t.e_b0 = __shfl_sync(0xffffffffu, ts[j].e_b0, src); t.e_c1 = __shfl_sync(0xffffffffu, ts[j].e_c1, src);
            t.base_inv = __shfl_sync(0xffffffffu, ts[j].base_inv, src);
            t.z0 = __shfl_sync(0xffffffffu, ts[j].z0, src); t.z1 = __shfl_sync(0xffffffffu, ts[j].z1, src);
            t.z2 = __shfl_sync(0xffffffffu, ts[j].z2, src);
            const int x0 = __shfl_sync(0xffffffffu, rect[j].x0, src), x1 = __shfl_sync(0xffffffffu, rect[j].x1, src);
            const int r0 = __shfl_sync(0xffffffffu, rect[j].r0, src);
            const int n = __shfl_sync(0xffffffffu, cells[j], src);
            const uint32_t tid_ = __shfl_sync(0xffffffffu, id[j], src);
            const int w = x1 - x0 + 1;
            for (int c0 = 0; c0 < n; c0 += 32) {
                const int c = c0 + lane;
                float beta = 0.f, gamma = 0.f;
                bool in = false;
                uint32_t pixel = 0;
                if (c < n) {
                    const int row = c / w, px = x0 + (c - row * w), py = r0 + row;
                    in = inside_test(t, (float)px, (float)py, beta, gamma);
                    pixel = (uint32_t)((H - 1 - py) * W + px);
                }
                count = append(q, count, lane, in, beta, gamma, t, pixel, tid_, keys);
            }
        }
    }
    drain(q, count, lane, keys);
}

// One WARP per deferred triangle: the same arithmetic, pixels of the bbox spread over the lanes.
__global__ void __launch_bounds__(kBigThreads) raster_big_kernel(FullRasterArgs a) {
    const int n_big = min(*a.big_count, a.big_cap);
    const int lane = threadIdx.x & 31, warps = gridDim.x * (kBigThreads / 32);
    for (int e = blockIdx.x * (kBigThreads / 32) + (threadIdx.x >> 5); e < n_big; e += warps) {
        const int2 item = a.big_list[e];
        const int pose_i = item.x, t = item.y & ((1 << 26) - 1);
        const uint32_t perm = (uint32_t)item.y >> 26;
        const Mat4 pose = a.poses[pose_i];
        const int W = a.cam.width, H = a.cam.height;
        const float wf = (float)W, hf = (float)H;
        const float4* rec = reinterpret_cast<const float4*>(a.models.cl_recs + (int64_t)(t >> 1) * kQuadFloats);
        const float4 r0 = __ldg(rec), r1 = __ldg(rec + 1), r2 = __ldg(rec + 2);
        ScreenVertex v[4];
        v[0] = transform_vertex(pose, a.cam.proj, wf, hf, r0.x, r0.y, r0.z);
        v[1] = transform_vertex(pose, a.cam.proj, wf, hf, r0.w, r1.x, r1.y);
        v[2] = transform_vertex(pose, a.cam.proj, wf, hf, r1.z, r1.w, r2.x);
        v[3] = transform_vertex(pose, a.cam.proj, wf, hf, r2.y, r2.z, r2.w);
        const ScreenVertex A = pick(v, perm & 3u), B = pick(v, (perm >> 2) & 3u), C = pick(v, perm >> 4);
        float bminx, bmaxx, bminy, bmaxy;
        bbox_of(A, B, C, (float)(W - 1), (float)(H - 1), bminx, bmaxx, bminy, bmaxy);
        const PixelRect r = pixel_rect_of(bminx, bmaxx, bminy, bmaxy, W, H);
        if (r.x0 > r.x1 || r.r0 > r.r1) continue;
        TriSetup ts;
        setup_edges(A, B, C, ts);
        const uint32_t id = a.key_pose_id ? (uint32_t)pose_i : (uint32_t)__ldg(a.models.cl_orig + t);
        unsigned long long* keys = a.keys + (a.single ? 0 : (size_t)pose_i * W * H);
        const int w = r.x1 - r.x0 + 1;
        const int cells = w * (r.r1 - r.r0 + 1);  // <= W*H
        for (int c = lane; c < cells; c += 32) {
            const int p1 = r.r0 + c / w, p0 = r.x0 + c % w;
            int32_t d;
            if (fragment_depth(ts, (float)p0, (float)p1, d)) atomicMin(keys + (size_t)(H - 1 - p1) * W + p0, make_key(d, id));
        }
    }
}

struct TreeAcc {
    int flags = 0, clutter = 0, total = 0;
};

// One covered pixel: key -> (depth, rgb) with the source-occlusion rule, and the tree statistics of the pose.
template <bool TREE, bool RGB>
__device__ __forceinline__ void resolve_pixel(const FullRasterArgs& a, unsigned long long key, int i, bool use_label,
                                              int pose_label, int32_t& d, uint32_t& rgb, TreeAcc& acc) {
    d = (int32_t)((uint32_t)(key >> 32) ^ 0x80000000u);
    const int32_t src = __ldg(a.scene.depth + i);
    const int src_label = (use_label && a.scene.label) ? (int)__ldg(a.scene.label + i) : 0;
    if (TREE) {
        // image_renderer.cuh:163-206 on the pixel's minimum depth (see pose_kernels.cu / DESIGN 4.11)
        ++acc.total;
        const int ad = abs(d - src);
        const bool gate = use_label ? (pose_label != src_label - 1 && ad >= 1) : ((float)ad > a.occlusion_threshold);
        if (gate && src > 0) {
            if (d > src) { acc.flags |= 2; ++acc.clutter; }
            else acc.flags |= 1;
        }
    }
    rgb = 0;
    if (source_occludes(d, src, use_label, pose_label, src_label, a.occlusion_threshold)) d = 0;
    else if (RGB) rgb = __ldg(a.models.tri_color + (uint32_t)key);
    if (d == INT_MAX) d = 0;  // max2zero (image_renderer.cuh:324-333,465-466)
}

// keys -> images.  grid.y = pose of the chunk; one thread per group of four consecutive pixels (no loop).  The common
// case -- four empty keys -- is two 16-byte loads, a compare and the zero stores.  Every key read is left empty again.
template <bool TREE, bool RGB>
__global__ void __launch_bounds__(256) resolve_full_kernel(FullRasterArgs a) {
    const int npx = a.cam.width * a.cam.height;
    const int img = blockIdx.y;
    const int i0 = 4 * (blockIdx.x * 256 + threadIdx.x);
    TreeAcc acc;
    if (i0 < npx) {
        unsigned long long* keys = a.keys + (size_t)img * npx + i0;
        unsigned long long key[4] = {~0ull, ~0ull, ~0ull, ~0ull};
        const bool full = i0 + 3 < npx;
        // keys outside the rectangle the rasterising kernels touched for this pose are empty: not even read
        bool touched = true;
        if (a.rect_min) {
            const int W = a.cam.width;
            const int y0 = i0 / W, x0 = i0 - y0 * W, y1 = (i0 + 3) / W, x1 = x0 + 3;
            const int rx0 = a.rect_min[2 * img], ry0 = a.rect_min[2 * img + 1], rx1 = a.rect_max[2 * img], ry1 = a.rect_max[2 * img + 1];
            // a group that wraps into the next row (W % 4 != 0) is treated as spanning both rows and every column
            touched = (y1 >= ry0 && y0 <= ry1) && (y0 != y1 || (x1 >= rx0 && x0 <= rx1));
        }
        if (!touched) {
        } else if (full && a.wide_stores) {  // npx % 4 == 0: 32-byte aligned group
            const ulonglong2 k01 = *reinterpret_cast<const ulonglong2*>(keys);
            const ulonglong2 k23 = *reinterpret_cast<const ulonglong2*>(keys + 2);
            key[0] = k01.x; key[1] = k01.y; key[2] = k23.x; key[3] = k23.y;
        } else {
            for (int k = 0; k < 4 && i0 + k < npx; ++k) key[k] = keys[k];
        }
        int32_t d4[4] = {0, 0, 0, 0};
        uint32_t r4 = 0, g4 = 0, b4 = 0;
        if ((key[0] & key[1] & key[2] & key[3]) != ~0ull) {
            const bool use_label = a.pose_label != nullptr;
            const int pose_label = use_label ? a.pose_label[img] : 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                if (key[k] == ~0ull) continue;
                keys[k] = ~0ull;
                uint32_t rgb;
                resolve_pixel<TREE, RGB>(a, key[k], i0 + k, use_label, pose_label, d4[k], rgb, acc);
                r4 |= (rgb & 0xFFu) << (8 * k);
                g4 |= ((rgb >> 8) & 0xFFu) << (8 * k);
                b4 |= ((rgb >> 16) & 0xFFu) << (8 * k);
            }
        }
        if (a.img_depth) {
            const size_t o = (size_t)img * npx + (size_t)i0;
            if (full && a.wide_stores) {
                *reinterpret_cast<int4*>(a.img_depth + o) = make_int4(d4[0], d4[1], d4[2], d4[3]);
                if (RGB) {
                    *reinterpret_cast<uint32_t*>(a.img_r + o) = r4;
                    *reinterpret_cast<uint32_t*>(a.img_g + o) = g4;
                    *reinterpret_cast<uint32_t*>(a.img_b + o) = b4;
                }
            } else {
                for (int k = 0; k < 4 && i0 + k < npx; ++k) {
                    a.img_depth[o + k] = d4[k];
                    if (RGB) {
                        a.img_r[o + k] = (uint8_t)(r4 >> (8 * k)); a.img_g[o + k] = (uint8_t)(g4 >> (8 * k));
                        a.img_b[o + k] = (uint8_t)(b4 >> (8 * k));
                    }
                }
            }
        }
    }
    if (TREE) {
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) {
            acc.flags |= __shfl_xor_sync(0xffffffffu, acc.flags, o);
            acc.clutter += __shfl_xor_sync(0xffffffffu, acc.clutter, o);
            acc.total += __shfl_xor_sync(0xffffffffu, acc.total, o);
        }
        if ((threadIdx.x & 31) == 0) {
            if (acc.flags) atomicOr(a.tree_acc + 3 * img, acc.flags);
            if (acc.clutter) atomicAdd(a.tree_acc + 3 * img + 1, acc.clutter);
            if (acc.total) atomicAdd(a.tree_acc + 3 * img + 2, acc.total);
        }
    }
}

__global__ void tree_finish_kernel(FullRasterArgs a) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= a.n_poses) return;
    const int flags = a.tree_acc[3 * p];
    const float clutter = (float)a.tree_acc[3 * p + 1], total = (float)a.tree_acc[3 * p + 2];
    if (a.tree_occluded) a.tree_occluded[p] = flags & 1;
    if (a.tree_occluded_other) a.tree_occluded_other[p] = (flags >> 1) & 1;
    // thrust::divides then multiplies by 100 (image_renderer.cuh:467-482); 0/0 = NaN like the reference
    if (a.tree_clutter) a.tree_clutter[p] = __fmul_rn(__fdiv_rn(clutter, total), 100.f);
    a.tree_total[p] = total;
}

}  // namespace

cudaError_t launch_full_raster(const FullRasterArgs& a, int max_model_quads, cudaStream_t stream) {
    if (a.n_poses <= 0 || max_model_quads <= 0) return cudaSuccess;
    cudaError_t e = cudaMemsetAsync(a.big_count, 0, sizeof(int), stream);
    if (e != cudaSuccess) return e;
    dim3 grid((max_model_quads + kFullThreads - 1) / kFullThreads, a.n_poses);
    raster_quads_kernel<<<grid, kFullThreads, 0, stream>>>(a);
    raster_big_kernel<<<a.big_ctas, kBigThreads, 0, stream>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_full_resolve(const FullRasterArgs& a, int n_images, cudaStream_t stream) {
    if (n_images <= 0) return cudaSuccess;
    const int npx = a.cam.width * a.cam.height;
    if (a.tree_total) {
        cudaError_t e = cudaMemsetAsync(a.tree_acc, 0, 3 * sizeof(int) * (size_t)n_images, stream);
        if (e != cudaSuccess) return e;
    }
    dim3 grid(((npx + 3) / 4 + 255) / 256, n_images);
    const bool rgb = a.img_depth && a.img_r;
    if (a.tree_total) {
        if (rgb) resolve_full_kernel<true, true><<<grid, 256, 0, stream>>>(a);
        else resolve_full_kernel<true, false><<<grid, 256, 0, stream>>>(a);
    } else {
        if (rgb) resolve_full_kernel<false, true><<<grid, 256, 0, stream>>>(a);
        else resolve_full_kernel<false, false><<<grid, 256, 0, stream>>>(a);
    }
    if (a.tree_total) tree_finish_kernel<<<(a.n_poses + 127) / 128, 128, 0, stream>>>(a);
    return cudaGetLastError();
}

namespace {

int env_or(const char* name, int dflt) {
    const char* v = getenv(name);
    return (v && *v) ? atoi(v) : dflt;
}

constexpr int kBigCap = 1 << 20;

}  // namespace

cudaError_t full_raster_job(::perch_ctx* c, const FullJob& j, int* launches) {
    if (j.n_poses <= 0) return cudaSuccess;
    cudaError_t e;
    const size_t npx = (size_t)c->cam.width * c->cam.height;
    if (!c->d_full_big) {
        if ((e = cudaMalloc(&c->d_full_big, sizeof(int2) * (size_t)kBigCap)) != cudaSuccess) return e;
        if ((e = cudaMalloc(&c->d_full_count, sizeof(int))) != cudaSuccess) return e;
    }
    FullRasterArgs a;
    a.models = c->models; a.cam = c->cam; a.scene = c->scene;
    a.occlusion_threshold = j.occlusion_threshold;
    a.key_pose_id = j.key_pose_id;
    a.big_list = c->d_full_big; a.big_count = c->d_full_count; a.big_cap = kBigCap;
    a.big_ctas = 4 * c->sm_count;
    a.big_cells = std::max(env_or("PERCH_FULL_BIG_CELLS", 4096), 1);
    a.lane_cells = std::max(env_or("PERCH_FULL_LANE_CELLS", 16), 0);
    if (j.single_keys) {
        a.poses = j.poses; a.pose_model_map = j.pose_model_map; a.pose_label = j.pose_label; a.n_poses = j.n_poses;
        a.keys = j.single_keys; a.single = 1;
        if (launches) *launches += 2;
        return launch_full_raster(a, c->max_model_quads, c->stream);
    }
    // Poses go through in chunks whose key images fit the budget.  Measured on 1024 poses of the C2 mesh
    // (tools/render_probe.py): 32 MB 3.1 ms, 64 MB 2.3 ms, 128 MB 1.9 ms, 256 MB 1.7 ms -- few large launches beat L2
    // residency of the keys.  (Running the DRAM-bound resolve of chunk i on a second, high-priority stream underneath
    // the issue-bound rasterising of chunk i + 1 was worth 2-4 % and is not kept.)
    const size_t budget = (size_t)std::max(env_or("PERCH_FULL_CHUNK_MB", 256), 1) << 20;
    const size_t chunk = std::min<size_t>({std::max<size_t>(budget / (npx * 8), 1), (size_t)j.n_poses, (size_t)65535});  // grid.y
    if (c->full_keys_n < chunk * npx) {
        if (c->d_full_keys) cudaFree(c->d_full_keys);
        c->d_full_keys = nullptr; c->full_keys_n = 0;
        if ((e = cudaMalloc(&c->d_full_keys, chunk * npx * 8)) != cudaSuccess) return e;
        if ((e = cudaMemsetAsync(c->d_full_keys, 0xFF, chunk * npx * 8, c->stream)) != cudaSuccess) return e;
        c->full_keys_n = chunk * npx;
    }
    if (j.tree_total && c->full_acc_n < 3 * (size_t)j.n_poses) {
        if (c->d_full_acc) cudaFree(c->d_full_acc);
        c->d_full_acc = nullptr; c->full_acc_n = 0;
        if ((e = cudaMalloc(&c->d_full_acc, 3 * sizeof(int) * (size_t)j.n_poses)) != cudaSuccess) return e;
        c->full_acc_n = 3 * (size_t)j.n_poses;
    }
    a.keys = c->d_full_keys;
    if (c->full_rects_n < 4 * chunk) {
        if (c->d_full_rects) cudaFree(c->d_full_rects);
        c->d_full_rects = nullptr; c->full_rects_n = 0;
        if ((e = cudaMalloc(&c->d_full_rects, 4 * sizeof(int) * chunk)) != cudaSuccess) return e;
        c->full_rects_n = 4 * chunk;
    }
    a.rect_min = c->d_full_rects;
    a.rect_max = c->d_full_rects + 2 * chunk;
    auto aligned = [](const void* p, uintptr_t m) { return (reinterpret_cast<uintptr_t>(p) & (m - 1)) == 0; };
    a.wide_stores = npx % 4 == 0 && aligned(j.img_depth, 16) && aligned(j.img_r, 4) && aligned(j.img_g, 4) && aligned(j.img_b, 4);
    for (size_t p0 = 0; p0 < (size_t)j.n_poses; p0 += chunk) {
        const int n = (int)std::min(chunk, (size_t)j.n_poses - p0);
        a.poses = j.poses + p0; a.pose_model_map = j.pose_model_map + p0;
        a.pose_label = j.pose_label ? j.pose_label + p0 : nullptr;
        a.n_poses = n;
        a.img_depth = j.img_depth ? j.img_depth + p0 * npx : nullptr;
        a.img_r = j.img_r ? j.img_r + p0 * npx : nullptr;
        a.img_g = j.img_g ? j.img_g + p0 * npx : nullptr;
        a.img_b = j.img_b ? j.img_b + p0 * npx : nullptr;
        a.tree_occluded = j.tree_occluded ? j.tree_occluded + p0 : nullptr;
        a.tree_occluded_other = j.tree_occluded_other ? j.tree_occluded_other + p0 : nullptr;
        a.tree_clutter = j.tree_clutter ? j.tree_clutter + p0 : nullptr;
        a.tree_total = j.tree_total ? j.tree_total + p0 : nullptr;
        a.tree_acc = j.tree_total ? c->d_full_acc + 3 * p0 : nullptr;
        if ((e = cudaMemsetAsync(a.rect_min, 0x7f, 2 * sizeof(int) * chunk, c->stream)) != cudaSuccess) return e;
        if ((e = cudaMemsetAsync(a.rect_max, 0x80, 2 * sizeof(int) * chunk, c->stream)) != cudaSuccess) return e;
        if ((e = launch_full_raster(a, c->max_model_quads, c->stream)) == cudaSuccess) e = launch_full_resolve(a, n, c->stream);
        if (e != cudaSuccess) {
            cudaFree(c->d_full_keys);  // the keys may be dirty: start from a fresh, cleared buffer next time
            c->d_full_keys = nullptr; c->full_keys_n = 0;
            return e;
        }
        if (launches) *launches += 3 + (j.tree_total ? 1 : 0);
    }
    return cudaSuccess;
}

}  // namespace perch
