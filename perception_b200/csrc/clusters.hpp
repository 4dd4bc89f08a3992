// Host-side mesh preprocessing for the candidate test of the per-pose kernels (pose_kernels.cu, pass A).
//
// The reference hands over a triangle soup (3 x float3 per triangle, cuda_renderer/src/model.cpp:60-135) and tests
// every triangle against every pose.  With a stride lattice (compute_point_clouds.cuh:47-56) the vast majority of
// triangles of a fine mesh cover no lattice pixel, so the kernels cull hierarchically:
//   * triangles that share an edge (two vertices with identical bit patterns) are paired into QUADS of four distinct
//     vertices, so that the per-triangle candidate test transforms four vertices for two triangles instead of six;
//   * two neighbouring quads form a CLUSTER (kClusterTris = 4 triangle slots) with a bounding sphere.  A cluster whose
//     projected sphere contains no lattice pixel is rejected with one test; only quads of surviving clusters get the
//     per-triangle test, and only its survivors the exact rasteriser.
// Exactness is untouched: culling is conservative, every survivor is rasterised with the reference's float sequence
// from the original coordinates IN THE ORIGINAL VERTEX ORDER (the `perm` field), and the colour tie-break uses the
// ORIGINAL triangle index (`orig`).
#pragma once
#include <cstdint>
#include <vector>

namespace perch {

constexpr int kClusterTris = 4;                      // triangle slots per cluster: slot j belongs to quad j >> 1
constexpr int kQuadVerts = 4;
constexpr int kClusterQuads = kClusterTris / 2;
constexpr int kQuadFloats = 3 * kQuadVerts;                      // one quad: v0.xyz v1.xyz v2.xyz v3.xyz = three float4
constexpr int kClusterFloats = kQuadFloats * kClusterQuads;      // record: 2 quads, 96 bytes
constexpr uint32_t kPermInvalid = 63;                // slot that repeats its quad's other triangle (never a candidate)

struct ClusterSet {
    // n_clusters x kClusterFloats: rec[c][q][3 * i + a] = coordinate a of vertex i of quad q (a quad is read with three
    // 16-byte loads).  Vertices 0 and 1 are the shared edge, vertex 2 belongs to the quad's first triangle only, vertex 3
    // to the second only; a quad holding a single triangle has vertex 3 = vertex 2.
    std::vector<float> recs;
    std::vector<float> bounds;   // n_clusters x 4: centre xyz, radius (inf when a coordinate is not finite)
    // n_clusters x kClusterQuads x 4: bounding sphere of every QUAD (centre xyz, radius rounded up; inf when a
    // coordinate is not finite; a slot that holds no triangle of its own repeats quad 0) -- level 1 of the candidate test
    std::vector<float> qbounds;
    std::vector<int> orig;       // n_clusters x kClusterTris: original (global) triangle index of every slot
    // n_clusters: 6 bits per slot, (i0 | i1 << 2 | i2 << 4) = quad vertices of the triangle's v0, v1, v2 in the
    // ORIGINAL order, or kPermInvalid for a slot without a triangle of its own
    std::vector<uint32_t> perm;
    std::vector<int> model_low, model_high;  // cluster range [low, high) per model
};

// tri9: n_tris x 9 floats (v0 v1 v2) in the original order; [low[m], high[m]) = triangles of model m.
ClusterSet build_clusters(const float* tri9, const std::vector<int>& low, const std::vector<int>& high);

}  // namespace perch
