// kernels.h -- host-callable launchers of the sm_100a kernels (implemented in *.cu).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include <vector>

namespace b200 {

// shade-stage records: float4s per triangle in the attribute pool ((vN0,u0) (vN1,u1) (vN2,u2) (v0,v1,v2,LOD) (T,alpha0) (B,alpha1)
// (alpha2, ltriIdx + 1 as integer bits, -,-)) and per material in the material table (wavefront.cu::shadeKernel: 6 base float4, then two per extra map)
constexpr int TRI_ATTR_STRIDE = 7, MAT_STRIDE = 14;

// Arguments of one extend (closest-hit) or shadow (any-hit) wave.
struct TraceArgs {
	const uint4* nodes;        // wide node pool, 5 x uint4 per node (BLAS nodes of all meshes, then the TLAS)
	const float4* tris;        // 3 x float4 per triangle, leaf order
	const float4* instances;   // 6 x float4 per instance record, TLAS leaf order
	const float4* rayO;        // (Ox,Oy,Oz,tmin)
	const float4* rayD;        // (Dx,Dy,Dz,tmax)
	const float4* hostO;       // optional zero-copy source: ray planes in pinned host memory, read over PCIe at ray fetch and written
	const float4* hostD;       //   through to stageO/stageD (== rayO/rayD, device), from where a lane re-reads its world-space ray
	float4* stageO;
	float4* stageD;
	float4* hits;              // closest-hit output, 16 B/ray           (extend)
	uint32_t* occluded;        // 1 bit/ray, pre-cleared, bit set = hit  (shadow); may be null when `accumulator` is set
	const float4* potentials;  // optional fused connect (shadow): E4 = (rgb, pixelIdx bits) per shadow ray ...
	float4* accumulator;       // ... added to accumulator[pixelIdx] when the ray is NOT occluded (.optix.cu:137-149, connections.h:23-34)
	uint32_t* fetchCounter;    // persistent-thread ray fetch counter (zeroed by the launcher, unless fetchCounterNext is set)
	uint32_t* fetchCounterNext;// optional: the counter of the NEXT wave on this stream; the kernel zeroes it, the launcher then issues no memset
	                           // (wave chains of a frame alternate between two counters that are zero when the frame starts)
	const uint32_t* countPtr;  // optional device-resident ray count (wavefront counters); overrides `count`
	uint32_t count;
	uint32_t tlasRoot;         // absolute node index of the top-level root
	uint32_t singleInstance;   // 1: the scene is one instance (record 0): rays enter it directly, the top-level node is not visited
	float tlasTScale;          // top-level culling: hit distances are kept in the hit instance's OBJECT-space units (as the reference does); a world-space box
	                           // entered at distance d can only hold hits of object-space distance >= d / tlasTScale (1 for rigid instances), see UpdateToplevel
	uint32_t magic;            // byte-decode constant, set by the launchers; passed as an argument on purpose, see qbyte() / qpair() in traverse.cu
	uint32_t* stepNodes;       // optional per-ray counters (COUNT variants)
	uint32_t* stepTris;
};

enum IsectMode { ISECT_REFERENCE = 0, ISECT_WATERTIGHT = 1 };

// grid = persistent CTAs (multiple of the SM count); returns cudaGetLastError()
cudaError_t LaunchExtend( const TraceArgs& a, IsectMode mode, bool countSteps, int smCount, cudaStream_t s );
cudaError_t LaunchShadow( const TraceArgs& a, IsectMode mode, int smCount, cudaStream_t s );

// The extend wave over host buffers with the kernel's scratch (fetch counter, device staging of rays and hits) in a file-scope device
// array of fixed size: no cudaMalloc / cudaFree anywhere on the path (traverse.cu).  n <= NOALLOC_MAX_RAYS; synchronises `s` before returning.
constexpr uint32_t NOALLOC_MAX_RAYS = 65536; // 256 B + 48 B per ray = 3 MB of device memory per GPU that loads the library
cudaError_t TraceClosestNoAlloc( const TraceArgs& base, IsectMode mode, int smCount, int device, const float* hostO4, const float* hostD4, uint32_t n, float* hostHits,
	cudaStream_t s );

// ---- refit (refit.cu) -------------------------------------------------------------------------------
struct RefitArgs {
	uint4* nodes;              // this mesh's nodes (absolute pool pointer)
	float4* tris;              // pool
	const float4* verts;       // new positions, float4[3T], ORIGINAL primitive order
	float4* nodeLo; float4* nodeHi; // scratch: float AABB per node of this mesh
	float4* shade;             // pool: (geometric normal, material bits) per ORIGINAL primitive of every mesh; normals are refreshed
	uint32_t nodeBase;         // absolute index of the mesh root in the pool
	uint32_t triBase, triCount;
	const uint2* leafInfo;     // per leaf-order triangle of this mesh: (original primitive, the shading record's .w), LaunchRefitLeafInfo; null: read from the records
};
cudaError_t LaunchRefitTop( const RefitArgs& a, const uint32_t* levelStart, int nTop, cudaStream_t s ); // levels [0,nTop) in one launch
cudaError_t LaunchRefitLevel( const RefitArgs& a, uint32_t levelBegin, uint32_t levelEnd, cudaStream_t s );
// the whole tree in one launch (refit.cu::refitWalkKernel).  parentInfo / arrived / stashOf: one word per node of the mesh each, filled once per
// topology by LaunchRefitParents (nodes = pool pointer, nodeBase = the mesh's root), which also leaves in *stashCount (device) the number of
// nodes with inner children; stash = 16 float4 (the eight slot boxes) for each of those
cudaError_t LaunchRefitParents( const uint4* nodes, uint32_t nodeBase, uint32_t nodeCount, uint32_t* parentInfo, uint32_t* arrived, uint32_t* stashOf, uint32_t* stashCount, cudaStream_t s );
cudaError_t LaunchRefitLeafInfo( const float4* tris, const float4* shade, uint32_t triBase, uint32_t triCount, uint2* out, cudaStream_t s );
cudaError_t LaunchRefitWalk( const RefitArgs& a, uint32_t nodeCount, const uint32_t* parentInfo, uint32_t* arrived, const uint32_t* stashOf, float4* stash, cudaStream_t s );

// ---- GPU build of the wide-BVH topology (build.cu); the caller runs the refit afterwards -----------------------
cudaError_t GpuBuildTopology( const float4* dVerts, uint32_t T, uint4* nodes, float4* tris, uint32_t nodeBase, uint32_t triBase, uint32_t maxNodes,
	cudaStream_t s, std::vector<uint32_t>& levelStart, uint32_t& nodeCount, int builder /* 1: Morton radix tree (LBVH), 2: agglomerative clustering along the curve (PLOC), 3: SAH inside the radix tree's cells, 4: binned SAH from the top */,
	void** keptScratch = nullptr, size_t* keptBytes = nullptr /* scratch owned by the caller and reused between builds (grown as needed); null: allocated and freed inside */ );

// ---- top-level build on the device (build.cu::tlasBuildKernel) -----------------------------------------------------------------
// what the host hands over per instance (in instance-id order); the kernel derives world boxes, tree, instance records and the
// shading-time instance table from it
struct TlasInstanceInput {
	float transform[16];      // row-major mat4 (common_types.h:469-632)
	float inv[16];            // mat4::Inverted of it, computed on the host with the reference's arithmetic (rendercore.cpp::Mat4Inverted)
	float lo[3]; uint32_t blasRoot;   // object-space bounds of the mesh, absolute node index of its root
	float hi[3]; uint32_t triBase;    //   ... first triangle of the mesh in the pools
	uint32_t meshId, flags, empty, pad; // flags: bit0 = inv is affine; empty: the mesh has no triangles (its box is the instance's origin)
};
static_assert( sizeof( TlasInstanceInput ) == 176, "TlasInstanceInput: eleven float4" );
constexpr uint32_t MAX_DEVICE_TLAS_INSTANCES = 2048; // one CTA builds the whole top level in shared memory; larger scenes use the host builder
// nodes: pool pointer of the top-level root; instRecs: 6 float4 per instance, leaf order; instShade: 4 float4 per instance id; info[0..3] = node count,
// level count, overflow flag, -.  scratch: at least TlasScratchBytes( n ) bytes.  One launch, asynchronous.
size_t TlasScratchBytes( uint32_t n );
cudaError_t LaunchTlasBuild( const TlasInstanceInput* dInstances, uint32_t n, uint4* nodes, uint32_t nodeBase, uint32_t maxNodes, float4* instRecs, float4* instShade,
	uint32_t* info, void* scratch, float primCost, cudaStream_t s );

// ---- SVGF-style filter + TAA (filter.cu; lib/CUDA/shared_kernel_code/finalize_shared.h:102-483) -------------------------------------------
// the per-pixel buffers of a filtered frame (device pointers, width * height elements; accumulator: two planes, direct then indirect)
struct FilterFrame {
	const float4* accumulator;
	uint4* features;             // x: albedo 10:11:11, y: packed normal (+ specular flag), z: depth, w: history length (4 bits) | specular flags << 4 | material << 6
	const float4* worldPos;      // xyz: first non-specular vertex, w: packed normal
	const float4* prevWorldPos;  // ... of the previous frame
	const float4* deltaDepth;    // zw: depth derivatives along x and y
	float4* shading;             // out: demodulated direct + indirect light, packed (see packLight)
	float2* motion;              // out: where the pixel was in the previous frame (+ 0.5)
	float4* moments;             // out: luminance moments (direct, direct^2, indirect, indirect^2)
	const float4* prevMoments;
	int width, height;
};
struct FilterReprojection {
	float4 prevPos, prevE, prevRight, prevUp; // the previous camera: position (w: plane offset), view direction, scaled screen axes (w: offsets)
	float j0, j1, prevj0, prevj1;             // TAA jitter of this and the previous frame
	float pixelValueScale, directClamp, indirectClamp;
	int camIsStationary;
};
FilterReprojection MakeFilterReprojection( const float* prevView /* pos, p1, p2, p3 */, int height, float j0, float j1, float prevj0, float prevj1, uint32_t spp,
	float directClamp, float indirectClamp, int camIsStationary );
cudaError_t LaunchFilterPrepare( const FilterFrame& f, const FilterReprojection& r, cudaStream_t s );
// A -> C, one a-trous pass (phase 1, 2, 3: steps 1, 2, 4); B = the previous frame's phase-1 output (read by phase 1 only)
cudaError_t LaunchFilterReconstruct( const FilterFrame& f, const float4* A, const float4* B, float4* C, int phase, int lastPass, cudaStream_t s );
cudaError_t LaunchFilterTaa( const float4* pixels, const float4* prevPixels, const float2* motion, float4* out, int w, int h, cudaStream_t s ); // out != pixels
cudaError_t LaunchFilterUnsharpen( const float4* pixels, float4* frame, int w, int h, cudaStream_t s );
cudaError_t LaunchFilterFinishNoTaa( const float4* pixels, float4* frame, int w, int h, cudaStream_t s );

} // namespace b200
