// Host-side mesh preprocessing for the candidate test of the per-pose kernels (pose_kernels.cu, pass A).
//
// The reference hands over a triangle soup (3 x float3 per triangle, cuda_renderer/src/model.cpp:60-135) and tests
// every triangle against every pose.  With a stride lattice (compute_point_clouds.cuh:47-56) the vast majority of
// triangles of a fine mesh cover no lattice pixel, so the kernels cull hierarchically: each model's triangles are put
// in Morton order of their centroids and cut into CLUSTERS of kClusterTris consecutive triangles with a bounding
// sphere.  A cluster whose projected sphere contains no lattice pixel is rejected with one test; only triangles of
// surviving clusters get the per-triangle candidate test, and only its survivors the exact rasteriser.
// Exactness is untouched: culling is conservative, every survivor is rasterised with the reference's float sequence
// from the original coordinates, and the colour tie-break uses the ORIGINAL triangle index (`orig`).
#pragma once
#include <cstdint>
#include <vector>

namespace perch {

constexpr int kClusterTris = 4;
constexpr int kClusterFloats = 9 * kClusterTris;  // record: 9 coordinate planes x kClusterTris lanes

struct ClusterSet {
    std::vector<float> recs;    // n_clusters x kClusterFloats: rec[c][k][j] = coordinate k (v0.x .. v2.z) of triangle j
    std::vector<float> bounds;  // n_clusters x 4: centre xyz, radius (inf when a coordinate is not finite)
    std::vector<int> orig;      // n_clusters x kClusterTris: original (global) triangle index; a short last cluster
                                // repeats its last triangle (an exact duplicate cannot change any result)
    std::vector<int> model_low, model_high;  // cluster range [low, high) per model
};

// tri9: n_tris x 9 floats (v0 v1 v2) in the original order; [low[m], high[m]) = triangles of model m.
ClusterSet build_clusters(const float* tri9, const std::vector<int>& low, const std::vector<int>& high);

}  // namespace perch
