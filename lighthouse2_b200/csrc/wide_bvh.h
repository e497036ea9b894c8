// wide_bvh.h -- acceleration-structure layout of RenderCore_B200 (host + device).
//
// Replaces, for the GPU, what BVH2::Rebuild (lib/RenderCore_Bart/bvh.cpp:57), optixAccelBuild
// (lib/rendercore_optix7/core_mesh.cpp:94-130) and rtpModelUpdate (RenderCore_OptixPrime_B/core_mesh.cpp:66)
// produce in the reference cores: an 8-wide BVH with child boxes quantised to 8 bits per plane
// relative to the node's own origin/power-of-two scale (80-byte nodes, five 16-byte loads), a
// triangle buffer in leaf order, and an instance table in top-level leaf order.
#pragma once
#include <algorithm>
#include <atomic>
#include <cstdint>
#include <string>
#include <thread>
#include <vector>

namespace b200 {

// ---- 80-byte wide node ---------------------------------------------------------------------------
// word layout (five uint4 loads n0..n4):
//  n0 = { px, py, pz, ex | ey<<8 | ez<<16 | imask<<24 }
//  n1 = { childBase, primBase, meta[0..3], meta[4..7] }
//  n2 = { qlox[0..3], qlox[4..7], qloy[0..3], qloy[4..7] }
//  n3 = { qloz[0..3], qloz[4..7], qhix[0..3], qhix[4..7] }
//  n4 = { qhiy[0..3], qhiy[4..7], qhiz[0..3], qhiz[4..7] }
// e* is a biased IEEE-754 exponent byte: plane = p + q * 2^(e-127).  imask bit s = slot s is an inner node.
// meta[s]: 0 = empty; inner = 0x20 | (24 + s); leaf = (unary prim count: 1->001, 2->011, 3->111) << 5 | offset
// of the leaf's first primitive from primBase (0..23).  Inner children sit contiguously at childBase in
// ascending slot order.  Indices are ABSOLUTE into the core's node / triangle / instance-reference pools.
struct WideNode {
	float p[3];
	uint8_t e[3];
	uint8_t imask;
	uint32_t childBase;
	uint32_t primBase;
	uint8_t meta[8];
	uint8_t qlox[8], qloy[8], qloz[8], qhix[8], qhiy[8], qhiz[8];
};
static_assert( sizeof( WideNode ) == 80, "WideNode must be five 16-byte words" );

// ---- 48-byte triangle record (leaf order) ---------------------------------------------------------
// v0 / e1 = v1-v0 / e2 = v2-v0 rounded exactly as Ray::IntersectTriangle computes them
// (lib/RenderCore_Bart/common.h:25-26), so storing the edges changes no bit of the reference arithmetic.
// v0.w carries the ORIGINAL primitive index (the id the caller sees) as integer bits.
struct WideTri { float v0[3]; uint32_t prim; float e1[3]; float pad1; float e2[3]; float pad2; };
static_assert( sizeof( WideTri ) == 48, "WideTri is three float4" );

// ---- 96-byte instance record (top-level leaf order) ------------------------------------------------
// inv = the full 4x4 inverse transform exactly as mat4::Inverted produces it (common_types.h:568-611);
// the reference transforms the ray with it at every TLAS leaf (bvh.cpp:607-610).
struct InstanceRec {
	float inv[16];
	uint32_t blasRoot;   // absolute node index of the mesh's root
	uint32_t instId;     // caller's instance index
	uint32_t meshId;
	uint32_t flags;      // bit0: inv is affine with last row exactly (0,0,0,1)
	float pad[4];
};
static_assert( sizeof( InstanceRec ) == 96, "InstanceRec is six float4" );

struct Aabb {
	float lo[3], hi[3];
	void reset() { lo[0] = lo[1] = lo[2] = 3.0e38f; hi[0] = hi[1] = hi[2] = -3.0e38f; }
	void grow( const Aabb& b ) { for (int a = 0; a < 3; a++) { if (b.lo[a] < lo[a]) lo[a] = b.lo[a]; if (b.hi[a] > hi[a]) hi[a] = b.hi[a]; } }
	void grow( const float* p ) { for (int a = 0; a < 3; a++) { if (p[a] < lo[a]) lo[a] = p[a]; if (p[a] > hi[a]) hi[a] = p[a]; } }
	float halfArea() const { float x = hi[0] - lo[0], y = hi[1] - lo[1], z = hi[2] - lo[2]; return x * y + y * z + z * x; }
};

// Per-lane traversal stack of the wave kernels (traverse.cu: STACK_SMEM + STACK_LOCAL entries, no bound check in the pass).  A visited
// bottom-level node pushes at most one entry (what is left of its inner-children group), a top-level node at most two (inner group +
// instance-leaf group), so a scene fits when blasLevels + 2 * tlasLevels <= TRAVERSAL_STACK_ENTRIES.  The core refuses deeper trees
// (SetGeometry / UpdateToplevel / the cache loader) instead of letting the kernels write past the stack.
constexpr int TRAVERSAL_STACK_ENTRIES = 48;
constexpr int MAX_TREE_LEVELS = TRAVERSAL_STACK_ENTRIES - 8; // one mesh's tree; leaves room for a top level of four levels

// Host loops over large meshes (builder passes, staging of triangle / shading records): fn( item ) for every item in [0, items), handed
// out one at a time to at most `threads` threads, the calling one included.  `perThread` = how many items make a thread worth starting.
template <class F> void ParallelFor( int threads, size_t items, F&& fn, size_t perThread = 1 )
{
	const size_t use = std::min<size_t>( threads > 1 ? (size_t)threads : 1, items / (perThread ? perThread : 1) );
	if (use <= 1) { for (size_t i = 0; i < items; i++) fn( i ); return; }
	std::atomic<size_t> next{ 0 };
	auto body = [&] { for (size_t i; (i = next.fetch_add( 1 )) < items;) fn( i ); };
	std::vector<std::thread> pool;
	for (size_t t = 1; t < use; t++) pool.emplace_back( body );
	body();
	for (std::thread& t : pool) t.join();
}
// the same over [0, n) in ranges of `grain`: fn( begin, end )
template <class F> void ParallelRange( int threads, size_t n, size_t grain, F&& fn )
{
	ParallelFor( threads, (n + grain - 1) / grain, [&]( size_t g ) { fn( g * grain, std::min( n, (g + 1) * grain ) ); }, 2 );
}

// Result of a host build (BLAS over triangles or TLAS over instance boxes).
struct WideBvh {
	std::vector<WideNode> nodes;         // breadth-first: level L occupies [levelStart[L], levelStart[L+1])
	std::vector<uint32_t> levelStart;    // size = levels + 1 (used by the level-by-level GPU refit)
	std::vector<uint32_t> primOrder;     // leaf-order slot -> original primitive index
	Aabb bounds;
};

extern int g_sahLevels;      // builder 3 on the device: how many levels of the radix tree's cells are rebuilt by SAH (Setting "gpuBuildLevels"; 1 = the bottoms only)
extern float g_sahPrimCost; // builder tunable: SAH cost of one leaf triangle relative to one wide-node visit

// Binned-SAH BVH2 over `n` primitive boxes, collapsed to the 8-wide layout above.
// maxLeafPrims: 3 for triangles (unary 3-bit count), 1 for instances.  Indices in the result are
// relative (root = 0, primBase relative to 0); the core rebases them when it places the tree in its pools.
void BuildWideBvh( const Aabb* primBoxes, uint32_t n, int maxLeafPrims, int threads, WideBvh& out );

// On-disk cache of built trees (bvh_cache.cpp): 64-bit content hash, load (false = absent or failed a check), store.
uint64_t HashBytes( const void* data, size_t bytes, uint64_t seed );
bool LoadWideBvh( const std::string& dir, uint64_t key, uint32_t triCount, int maxLeafPrims, WideBvh& out );
bool StoreWideBvh( const std::string& dir, uint64_t key, uint32_t triCount, int maxLeafPrims, const WideBvh& bvh );

// Re-quantise all child boxes of a wide tree from per-primitive boxes given in ORIGINAL index order
// (host reference implementation of the GPU refit; used by tests and for TLAS updates).
void RefitWideBvhHost( WideBvh& bvh, const Aabb* primBoxes );

// Quantisation of one node's frame: origin = lo, per-axis exponent byte with (hi-lo) <= 255 * 2^(e-127).
void QuantiseFrame( const Aabb& nodeBox, float p[3], uint8_t e[3] );
// Conservative 8-bit planes of a child box in a node frame (computed in double: floor/ceil with a
// 1/64-cell guard band so the float slab arithmetic on the device stays conservative).
void QuantiseChild( const float p[3], const uint8_t e[3], const Aabb& child, uint8_t qlo[3], uint8_t qhi[3] );

} // namespace b200
