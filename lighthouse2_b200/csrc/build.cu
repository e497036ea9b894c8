// build.cu -- GPU build of a mesh's 8-wide BVH topology.
//
// Replaces, for new or re-topologised meshes, the host SAH builder (bvh_builder.cpp) and with it what the reference does in
// BVH2::Rebuild (lib/RenderCore_Bart/bvh.cpp:57-256) / optixAccelBuild (lib/rendercore_optix7/core_mesh.cpp:94-130):
//   1. per-triangle boxes + centroid bounds                         (primBoundsKernel)
//   2. 63-bit Morton codes of the centroids, radix sort             (mortonKernel, cub::DeviceRadixSort -- a library sort, not a hot op)
//   3. binary radix tree over the sorted codes (Karras, "Maximizing Parallelism in the Construction of BVHs, Octrees, and
//      k-d Trees", HPG 2012)                                        (radixTreeKernel)
//   4. bottom-up boxes with one atomic counter per inner node      (binaryBoxesKernel)
//   5. top-down collapse to 8-wide nodes, one launch per level: a wide node starts from the two children of its binary node and
//      keeps opening the child with the largest area until it has eight; subtrees of at most three triangles become leaf
//      slots; children go to the octant slot they are farthest against (same rule as AssignSlots in bvh_builder.cpp); node and
//      triangle ranges are claimed with atomics, so a level is a contiguous index range as the refit expects (collapseKernel)
// Only the TOPOLOGY is written (imask, meta, childBase, primBase, the primitive tag of every leaf-order triangle slot): the boxes,
// frames, quantised planes and triangle records are then produced by the level-by-level refit (refit.cu), which a built mesh
// needs anyway.  Any conservative hierarchy gives the reference's closest hits (the BLAS traversal never culls by distance,
// bvh.cpp:281-298), so results equal the SAH tree's; the LBVH tree is faster to build and slower to trace.
#include "kernels.h"
#include "wide_bvh.h"
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#ifndef B200_DP_COLLAPSE
#define B200_DP_COLLAPSE 1 // children of a wide node from the SAH dynamic programme (1) or by opening the largest child (0)
#endif
#ifndef B200_GROUP_COLLAPSE
#define B200_GROUP_COLLAPSE 1 // 1: eight lanes per wide node (collapseNodeGroup); 0: one thread per node (collapseNode, the first version, kept for A/B)
#endif
#define B200_USE_GROUP_COLLAPSE (B200_GROUP_COLLAPSE && B200_DP_COLLAPSE)
#include <vector>
#include <chrono>
#include <stdio.h>
#include <stdlib.h>

namespace b200 {

namespace {

constexpr uint32_t LEAF = 0x80000000u; // child reference: bit 31 = sorted-leaf index, else inner binary node

__device__ __forceinline__ int orderedInt( const float f ) { const int i = __float_as_int( f ); return i >= 0 ? i : i ^ 0x7fffffff; }
__device__ __forceinline__ float orderedFloat( const int i ) { return __int_as_float( i >= 0 ? i : i ^ 0x7fffffff ); }

// 1. boxes of the triangles and the bounds of their centroids (ordered-int atomics; centroidBounds = {minx,miny,minz,maxx,maxy,maxz})
__global__ void primBoundsKernel( const float4* __restrict__ verts, const uint32_t T, float4* __restrict__ primLo, float4* __restrict__ primHi, int* centroidBounds )
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	float c[3] = { 0, 0, 0 };
	const bool valid = i < T;
	if (valid)
	{
		const float4 p0 = __ldg( verts + (size_t)i * 3 ), p1 = __ldg( verts + (size_t)i * 3 + 1 ), p2 = __ldg( verts + (size_t)i * 3 + 2 );
		const float lx = fminf( p0.x, fminf( p1.x, p2.x ) ), ly = fminf( p0.y, fminf( p1.y, p2.y ) ), lz = fminf( p0.z, fminf( p1.z, p2.z ) );
		const float hx = fmaxf( p0.x, fmaxf( p1.x, p2.x ) ), hy = fmaxf( p0.y, fmaxf( p1.y, p2.y ) ), hz = fmaxf( p0.z, fmaxf( p1.z, p2.z ) );
		primLo[i] = make_float4( lx, ly, lz, 0 ), primHi[i] = make_float4( hx, hy, hz, 0 );
		c[0] = 0.5f * (lx + hx), c[1] = 0.5f * (ly + hy), c[2] = 0.5f * (lz + hz);
	}
	for (int a = 0; a < 3; a++)
	{
		int lo = valid ? orderedInt( c[a] ) : 0x7fffffff, hi = valid ? orderedInt( c[a] ) : (int)0x80000000;
		lo = __reduce_min_sync( 0xffffffffu, lo ), hi = __reduce_max_sync( 0xffffffffu, hi );
		if ((threadIdx.x & 31) == 0) atomicMin( centroidBounds + a, lo ), atomicMax( centroidBounds + 3 + a, hi );
	}
}

__device__ __forceinline__ uint64_t spread21( uint64_t x )
{
	x &= 0x1fffffull;
	x = (x | x << 32) & 0x1f00000000ffffull;
	x = (x | x << 16) & 0x1f0000ff0000ffull;
	x = (x | x << 8) & 0x100f00f00f00f00full;
	x = (x | x << 4) & 0x10c30c30c30c30c3ull;
	x = (x | x << 2) & 0x1249249249249249ull;
	return x;
}

// 2. Morton code of every centroid on a 2^21 grid over the centroid bounds
__global__ void mortonKernel( const float4* __restrict__ primLo, const float4* __restrict__ primHi, const uint32_t T, const int* __restrict__ centroidBounds,
	uint64_t* __restrict__ keys, uint32_t* __restrict__ vals )
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= T) return;
	const float4 lo = primLo[i], hi = primHi[i];
	const float c[3] = { 0.5f * (lo.x + hi.x), 0.5f * (lo.y + hi.y), 0.5f * (lo.z + hi.z) };
	uint64_t q[3];
	for (int a = 0; a < 3; a++)
	{
		const float mn = orderedFloat( centroidBounds[a] ), mx = orderedFloat( centroidBounds[3 + a] );
		const float ext = mx - mn;
		float t = ext > 0 ? (c[a] - mn) / ext : 0.f;
		t = fminf( fmaxf( t, 0.f ), 1.f );
		uint32_t g = (uint32_t)(t * 2097151.0f);
		q[a] = g > 2097151u ? 2097151u : g;
	}
	keys[i] = spread21( q[0] ) << 2 | spread21( q[1] ) << 1 | spread21( q[2] );
	vals[i] = i;
}

// common-prefix length of sorted keys i and j (-1 outside the array); equal keys fall back to the index, which keeps the tree well formed
__device__ __forceinline__ int prefixLen( const uint64_t* __restrict__ keys, const int n, const int i, const int j )
{
	if (j < 0 || j >= n) return -1;
	const uint64_t a = keys[i], b = keys[j];
	return a != b ? __clzll( (long long)(a ^ b) ) : 64 + __clz( i ^ j );
}

// 3. Karras 2012, algorithm of section 4: inner node i of the binary radix tree over n sorted keys (n >= 2)
__global__ void radixTreeKernel( const uint64_t* __restrict__ keys, const int n, uint32_t* __restrict__ childL, uint32_t* __restrict__ childR,
	uint32_t* __restrict__ rangeFirst, uint32_t* __restrict__ rangeCount, uint32_t* __restrict__ parentInner, uint32_t* __restrict__ parentLeaf )
{
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n - 1) return;
	const int d = prefixLen( keys, n, i, i + 1 ) - prefixLen( keys, n, i, i - 1 ) >= 0 ? 1 : -1;
	const int dmin = prefixLen( keys, n, i, i - d );
	int lmax = 2;
	while (prefixLen( keys, n, i, i + lmax * d ) > dmin) lmax *= 2;
	int l = 0;
	for (int t = lmax / 2; t >= 1; t /= 2) if (prefixLen( keys, n, i, i + (l + t) * d ) > dmin) l += t;
	const int j = i + l * d;
	const int dnode = prefixLen( keys, n, i, j );
	int s = 0;
	for (int div = 2, t = (l + div - 1) / div; ; div *= 2, t = (l + div - 1) / div)
	{
		if (prefixLen( keys, n, i, i + (s + t) * d ) > dnode) s += t;
		if (t <= 1) break;
	}
	const int gamma = i + s * d + min( d, 0 );
	const int first = min( i, j ), last = max( i, j );
	const uint32_t L = first == gamma ? (LEAF | (uint32_t)gamma) : (uint32_t)gamma;
	const uint32_t R = last == gamma + 1 ? (LEAF | (uint32_t)(gamma + 1)) : (uint32_t)(gamma + 1);
	childL[i] = L, childR[i] = R, rangeFirst[i] = (uint32_t)first, rangeCount[i] = (uint32_t)(last - first + 1);
	if (L & LEAF) parentLeaf[gamma] = (uint32_t)i; else parentInner[gamma] = (uint32_t)i;
	if (R & LEAF) parentLeaf[gamma + 1] = (uint32_t)i; else parentInner[gamma + 1] = (uint32_t)i;
	if (i == 0) parentInner[0] = 0xffffffffu;
}

// 4. boxes of the binary inner nodes and the collapse plan: every sorted leaf walks up; the second thread to reach a node merges its
// children.  The plan is the dynamic programme of bvh_builder.cpp (after Ylitie, Karras, Laine, HPG 2017, sec. 4.1) run bottom-up:
// cost[i-1] = cheapest way to represent the subtree as a forest of at most i roots (i = 1: one wide node, or one leaf slot when it has
// at most three triangles), node cost = area, leaf cost = area * triangles * primCost.  planLo = cost[0..3], planHi = cost[4..6] and a
// word with the choices: bits 3(i-2)..3(i-2)+2 = left's share of i roots for i = 2..7 (0 = "use i-1 roots"), bits 18..20 = left's
// share of the 8 slots when the node becomes a wide node, bit 21 = a single root is a leaf slot.
__device__ __forceinline__ void loadPlan( const uint32_t ref, const uint32_t* __restrict__ sortedPrim, const float4* __restrict__ primLo, const float4* __restrict__ primHi,
	const float4* planLo, const float4* planHi, const float primCost, float c[7] )
{
	if (ref & LEAF)
	{
		const uint32_t p = sortedPrim[ref & ~LEAF];
		const float4 l = primLo[p], h = primHi[p];
		const float x = h.x - l.x, y = h.y - l.y, z = h.z - l.z, v = (x * y + y * z + z * x) * primCost;
		for (int i = 0; i < 7; i++) c[i] = v;
	}
	else
	{
		const float4 a = __ldcg( planLo + ref ), b = __ldcg( planHi + ref );
		c[0] = a.x, c[1] = a.y, c[2] = a.z, c[3] = a.w, c[4] = b.x, c[5] = b.y, c[6] = b.z;
	}
}

// the plan of one binary node from its two children's plans (see above).  maxLeaf: most primitives a leaf slot can hold (3 triangles, 1 instance)
__device__ __forceinline__ void mergePlan( const float cl[7], const float cr[7], const float area, const uint32_t count, const uint32_t maxLeaf, const float primCost,
	float cost[7], uint32_t& bits )
{
	bits = 0;
	float best8 = 3.0e38f;
	int k8 = 4;
	for (int k = 1; k < 8; k++) { const float v = cl[k - 1] + cr[8 - k - 1]; if (v < best8) best8 = v, k8 = k; }
	const float wideCost = area + best8, leafCost = count <= maxLeaf ? area * (float)count * primCost : 3.0e38f;
	cost[0] = fminf( leafCost, wideCost );
	bits |= (uint32_t)k8 << 18;
	if (leafCost <= wideCost) bits |= 1u << 21;
	for (int r = 2; r <= 7; r++)
	{
		float best = 3.0e38f;
		int bk = 1;
		for (int k = 1; k < r; k++) { const float v = cl[k - 1] + cr[r - k - 1]; if (v < best) best = v, bk = k; }
		if (best < cost[r - 2]) cost[r - 1] = best, bits |= (uint32_t)bk << (3 * (r - 2));
		else cost[r - 1] = cost[r - 2];
	}
}

__global__ void binaryBoxesKernel( const uint32_t* __restrict__ sortedPrim, const float4* __restrict__ primLo, const float4* __restrict__ primHi, const int n,
	const uint32_t* __restrict__ childL, const uint32_t* __restrict__ childR, const uint32_t* __restrict__ rangeCount, const uint32_t* __restrict__ parentInner,
	const uint32_t* __restrict__ parentLeaf, float4* nodeLo, float4* nodeHi, float4* planLo, float4* planHi, const float primCost, uint32_t* visits )
{
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	uint32_t node = parentLeaf[i];
	while (node != 0xffffffffu)
	{
		__threadfence();
		if (atomicAdd( visits + node, 1u ) == 0) return; // the sibling subtree is not done yet
		__threadfence();
		float4 lo[2], hi[2];
		const uint32_t c[2] = { childL[node], childR[node] };
		for (int k = 0; k < 2; k++)
		{
			if (c[k] & LEAF) { const uint32_t p = sortedPrim[c[k] & ~LEAF]; lo[k] = primLo[p], hi[k] = primHi[p]; }
			else lo[k] = __ldcg( nodeLo + c[k] ), hi[k] = __ldcg( nodeHi + c[k] );
		}
		const float4 nl = make_float4( fminf( lo[0].x, lo[1].x ), fminf( lo[0].y, lo[1].y ), fminf( lo[0].z, lo[1].z ), 0 );
		const float4 nh = make_float4( fmaxf( hi[0].x, hi[1].x ), fmaxf( hi[0].y, hi[1].y ), fmaxf( hi[0].z, hi[1].z ), 0 );
		__stcg( nodeLo + node, nl ), __stcg( nodeHi + node, nh );
		// collapse plan
		float cl[7], cr[7], cost[7];
		loadPlan( c[0], sortedPrim, primLo, primHi, planLo, planHi, primCost, cl );
		loadPlan( c[1], sortedPrim, primLo, primHi, planLo, planHi, primCost, cr );
		const float ex = nh.x - nl.x, ey = nh.y - nl.y, ez = nh.z - nl.z, area = ex * ey + ey * ez + ez * ex;
		uint32_t bits;
		mergePlan( cl, cr, area, rangeCount[node], 3, primCost, cost, bits );
		__stcg( planLo + node, make_float4( cost[0], cost[1], cost[2], cost[3] ) );
		__stcg( planHi + node, make_float4( cost[4], cost[5], cost[6], __uint_as_float( bits ) ) );
		node = parentInner[node];
	}
}

// ---- 3c. SAH inside the radix tree, level by level (builder 3).  The Morton splits are good at cutting a scene into compact cells and
// poor at deciding which triangles (or which cells) share a node: the order along the curve, not the surface area, decides.  The host
// builder's trees trace 4 % faster than the radix tree's; SAH in the bottom levels alone closes that for primary and diffuse rays, SAH
// above them as well for the long shadow rays (profiles/r02_experiments.md).  So the radix tree only supplies CELLS:
//   level 1: items = the triangles.  Every maximal radix subtree of at most SAHB_CLUSTER items is rebuilt from its root down by binned SAH
//            (16 bins on each axis; the split with the smallest sum of area * triangles over the three axes), one warp per subtree, the
//            items' boxes in shared memory.
//   level k + 1: items = the subtrees rebuilt so far (and what was left over); the same again, until the root is inside.
// Three levels at 1 M triangles.  A rebuilt subtree lives in the index places of the radix subtree it replaces: the radix subtree over the
// sorted positions [f, l] rooted at r uses the inner-node indices [f, l] without one end, r is the other end (Karras: the children of a
// split at g are g and g + 1).  Hence the rule that keeps the books at every level: an item over [f1, l1] with root r1 has the spare index
// (r1 == f1 ? l1 : f1) -- a single triangle at p has p --, a cluster of items may use its items' spare indices except its own root and
// its own spare end, and those are exactly as many as it needs.  Nothing moves in the sorted order; leaves stay LEAF | position.
// binaryBoxesKernel and the collapse then run as for the plain radix tree.
#ifndef SAHB_CLUSTER
#define SAHB_CLUSTER 256
#endif
#ifndef SAHB_BINS_N
#define SAHB_BINS_N 16
#endif
constexpr int SAHB_BINS = SAHB_BINS_N, SAHB_WARPS = SAHB_CLUSTER <= 256 ? 4 : (SAHB_CLUSTER <= 1024 ? 2 : 1); // (shared memory: 38 bytes per item and warp)

struct SahbSparse { uint32_t *flag, *ref, *last; };                         // per sorted position: an item starts here; its reference (LEAF | position, or its root) and last position
struct SahbItems { uint32_t *ref, *first, *last; float4 *lo, *hi; };       // the items of a level, in position order

__global__ void sahbInitKernel( const uint32_t n, const SahbSparse sp )
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) sp.flag[i] = 1, sp.ref[i] = LEAF | i, sp.last[i] = i;
}

// the level's item list from the flags and their running sum
__global__ void sahbCompactKernel( const uint32_t n, const SahbSparse sp, const uint32_t* __restrict__ before, const uint32_t* __restrict__ sortedPrim,
	const float4* __restrict__ primLo, const float4* __restrict__ primHi, const float4* __restrict__ nodeLo, const float4* __restrict__ nodeHi, const SahbItems it )
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n || !sp.flag[i]) return;
	const uint32_t k = before[i], ref = sp.ref[i];
	it.ref[k] = ref, it.first[k] = i, it.last[k] = sp.last[i];
	if (ref & LEAF) { const uint32_t p = sortedPrim[i]; it.lo[k] = primLo[p], it.hi[k] = primHi[p]; }
	else it.lo[k] = nodeLo[ref], it.hi[k] = nodeHi[ref];
}

// the roots of a level: radix nodes above the items with at most SAHB_CLUSTER items below them, whose parents have more
__global__ void sahbRootsKernel( const uint32_t n, const SahbSparse sp, const uint32_t* __restrict__ before, const SahbItems it, const uint32_t* __restrict__ rangeFirst,
	const uint32_t* __restrict__ rangeCount, const uint32_t* __restrict__ parentInner, uint4* __restrict__ roots, uint32_t* rootCount, const uint32_t limit )
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i + 1 >= n) return;
	// is inner node i still a node of the radix tree, or one of the indices a rebuilt subtree has taken?
	const uint32_t k = before[i] + sp.flag[i] - 1, f1 = it.first[k], l1 = it.last[k], r1 = it.ref[k];
	if (f1 != l1 && i != (r1 == f1 ? l1 : f1)) return;
	const uint32_t f = rangeFirst[i], l = f + rangeCount[i] - 1, items = before[l] + sp.flag[l] - before[f];
	if (items > limit) return;
	const uint32_t par = parentInner[i];
	if (par != 0xffffffffu)
	{
		const uint32_t pf = rangeFirst[par], pl = pf + rangeCount[par] - 1;
		if (before[pl] + sp.flag[pl] - before[pf] <= limit) return; // inside a larger subtree that is rebuilt as a whole
	}
	roots[atomicAdd( rootCount, 1u )] = make_uint4( i, before[f], items, l );
}

struct SahbWarp {
	float lo[3][SAHB_CLUSTER], hi[3][SAHB_CLUSTER];      // the items' boxes, where they were loaded: a split moves indices, not boxes
	uint32_t ref[SAHB_CLUSTER], weight[SAHB_CLUSTER];     // reference and triangles of the item
	uint32_t pool[SAHB_CLUSTER];                          // the items' spare indices
	uint16_t order[2][SAHB_CLUSTER];                      // the items in tree order: [0] current, [1] scratch of the partition
	int binLo[3][SAHB_BINS][3], binHi[3][SAHB_BINS][3], binWeight[3][SAHB_BINS], binItems[3][SAHB_BINS];
	uint32_t stackId[24], stackBegin[24], stackCount[24], stackWeight[24];
};

// CELLS: the subtrees are cells of the radix tree (items, spare indices: see above).  !CELLS (builder 4): the subtrees hang below the
// top-down phase -- roots[job] = (node, first position, triangles), the items are the triangles at those positions of `order`, and
// the subtree's other nodes are a block of fresh indices taken from *nodeCounter.
template <bool CELLS>
__global__ void __launch_bounds__( SAHB_WARPS * 32 ) sahbBuildKernel( const uint4* __restrict__ roots, const uint32_t* __restrict__ rootCount, const SahbItems it, const SahbSparse sparse,
	uint32_t* __restrict__ childL, uint32_t* __restrict__ childR, uint32_t* __restrict__ rangeCount, uint32_t* __restrict__ parentInner, uint32_t* __restrict__ parentLeaf,
	float4* __restrict__ nodeLo, float4* __restrict__ nodeHi,
	const uint32_t* __restrict__ order = nullptr, const float4* __restrict__ primLo = nullptr, const float4* __restrict__ primHi = nullptr, uint32_t* nodeCounter = nullptr )
{
	extern __shared__ __align__( 16 ) unsigned char sahbSmem[];
	SahbWarp& w = reinterpret_cast<SahbWarp*>( sahbSmem )[threadIdx.x >> 5];
	const uint32_t lane = threadIdx.x & 31, job = blockIdx.x * SAHB_WARPS + (threadIdx.x >> 5);
	if (job >= *rootCount) return; // (whole warps)
	const uint4 rt = roots[job];
	const uint32_t root = rt.x, item0 = rt.y, total = rt.z, firstPos = CELLS ? it.first[item0] : rt.y, lastPos = rt.w;
	const uint32_t spareEnd = !CELLS ? 0xffffffffu : (root == firstPos ? lastPos : firstPos);
	float blo[3] = { 3.0e38f, 3.0e38f, 3.0e38f }, bhi[3] = { -3.0e38f, -3.0e38f, -3.0e38f };
	uint32_t wsum = 0;
	uint32_t freshBase = 0;
	if (!CELLS)
	{
		if (lane == 0 && total > 2) freshBase = atomicAdd( nodeCounter, total - 2 );
		freshBase = __shfl_sync( 0xffffffffu, freshBase, 0 );
	}
	for (uint32_t j = lane; j < total; j += 32)
	{
		if (!CELLS)
		{
			const uint32_t p = order[firstPos + j];
			const float4 l = primLo[p], h = primHi[p];
			w.ref[j] = LEAF | (firstPos + j), w.weight[j] = 1, w.pool[j] = freshBase + j, w.order[0][j] = (uint16_t)j; // (total - 2 of them are used)
			w.lo[0][j] = l.x, w.lo[1][j] = l.y, w.lo[2][j] = l.z, w.hi[0][j] = h.x, w.hi[1][j] = h.y, w.hi[2][j] = h.z;
			wsum++;
			continue;
		}
		const uint32_t r = it.ref[item0 + j], f1 = it.first[item0 + j], l1 = it.last[item0 + j];
		const float4 l = it.lo[item0 + j], h = it.hi[item0 + j];
		w.ref[j] = r, w.weight[j] = l1 - f1 + 1, w.pool[j] = (r & LEAF) ? f1 : (r == f1 ? l1 : f1), w.order[0][j] = (uint16_t)j;
		w.lo[0][j] = l.x, w.lo[1][j] = l.y, w.lo[2][j] = l.z, w.hi[0][j] = h.x, w.hi[1][j] = h.y, w.hi[2][j] = h.z;
		blo[0] = fminf( blo[0], l.x ), blo[1] = fminf( blo[1], l.y ), blo[2] = fminf( blo[2], l.z ), bhi[0] = fmaxf( bhi[0], h.x ), bhi[1] = fmaxf( bhi[1], h.y ), bhi[2] = fmaxf( bhi[2], h.z );
		wsum += l1 - f1 + 1;
		sparse.flag[f1] = 0; // the item is inside from now on ...
	}
	for (int d = 16; d; d >>= 1)
	{
		for (int a = 0; a < 3; a++) blo[a] = fminf( blo[a], __shfl_xor_sync( 0xffffffffu, blo[a], d ) ), bhi[a] = fmaxf( bhi[a], __shfl_xor_sync( 0xffffffffu, bhi[a], d ) );
		wsum += __shfl_xor_sync( 0xffffffffu, wsum, d );
	}
	__syncwarp();
	if (CELLS && lane == 0)
	{
		// ... and the subtree is an item of the next level
		sparse.flag[firstPos] = 1, sparse.ref[firstPos] = root, sparse.last[firstPos] = lastPos;
		nodeLo[root] = make_float4( blo[0], blo[1], blo[2], 0 ), nodeHi[root] = make_float4( bhi[0], bhi[1], bhi[2], 0 );
	}
	uint32_t nextPool = 0, sp = 0;
	uint32_t id = root, begin = 0, count = total, weight = wsum;
	while (true)
	{
		// ---- one node: `count` >= 2 items at [begin, begin + count) of the tree order, `weight` triangles
		uint32_t nl = count >> 1, wl = 0xffffffffu; // the fallback: halves in curve order
		if (count > 2)
		{
			// centroid bounds (twice the centroid: lo + hi)
			float cmin[3] = { 3.0e38f, 3.0e38f, 3.0e38f }, cmax[3] = { -3.0e38f, -3.0e38f, -3.0e38f };
			for (uint32_t j = begin + lane; j < begin + count; j += 32)
			{
				const uint32_t o = w.order[0][j];
				for (int a = 0; a < 3; a++) { const float c = w.lo[a][o] + w.hi[a][o]; cmin[a] = fminf( cmin[a], c ), cmax[a] = fmaxf( cmax[a], c ); }
			}
			for (int a = 0; a < 3; a++) for (int d = 16; d; d >>= 1)
				cmin[a] = fminf( cmin[a], __shfl_xor_sync( 0xffffffffu, cmin[a], d ) ), cmax[a] = fmaxf( cmax[a], __shfl_xor_sync( 0xffffffffu, cmax[a], d ) );
			float scale[3];
			for (int a = 0; a < 3; a++) scale[a] = cmax[a] > cmin[a] ? (float)SAHB_BINS * 0.999f / (cmax[a] - cmin[a]) : 0.f;
			for (uint32_t t = lane; t < 3 * SAHB_BINS; t += 32)
			{
				int* bl = &w.binLo[0][0][0] + t * 3, *bh = &w.binHi[0][0][0] + t * 3;
				bl[0] = bl[1] = bl[2] = 0x7fffffff, bh[0] = bh[1] = bh[2] = (int)0x80000000, (&w.binWeight[0][0])[t] = 0, (&w.binItems[0][0])[t] = 0;
			}
			__syncwarp();
			for (uint32_t j = begin + lane; j < begin + count; j += 32)
			{
				int l[3], h[3];
				const uint32_t o = w.order[0][j];
				for (int a = 0; a < 3; a++) l[a] = orderedInt( w.lo[a][o] ), h[a] = orderedInt( w.hi[a][o] );
				const int wt = (int)w.weight[o];
				for (int a = 0; a < 3; a++)
				{
					const int b = max( 0, min( SAHB_BINS - 1, (int)((w.lo[a][o] + w.hi[a][o] - cmin[a]) * scale[a]) ) );
					for (int k = 0; k < 3; k++) atomicMin( &w.binLo[a][b][k], l[k] ), atomicMax( &w.binHi[a][b][k], h[k] );
					atomicAdd( &w.binWeight[a][b], wt ), atomicAdd( &w.binItems[a][b], 1 );
				}
			}
			__syncwarp();
			// candidates: axis a, the bins up to and including b go left (b < BINS - 1); 3 * 15 = 45 of them over the 32 lanes
			float bestCost = 3.0e38f;
			uint32_t bestCand = 0xffffffffu, bestLeft = 0, bestLeftWeight = 0;
			for (uint32_t t = lane; t < 3 * (SAHB_BINS - 1); t += 32)
			{
				const int a = t / (SAHB_BINS - 1), b = t % (SAHB_BINS - 1);
				if (scale[a] == 0.f) continue;
				float ll[3] = { 3.0e38f, 3.0e38f, 3.0e38f }, lh[3] = { -3.0e38f, -3.0e38f, -3.0e38f }, rl[3] = { 3.0e38f, 3.0e38f, 3.0e38f }, rh[3] = { -3.0e38f, -3.0e38f, -3.0e38f };
				int cl = 0, cr = 0, il = 0, ir = 0;
				for (int k = 0; k < SAHB_BINS; k++)
				{
					const int c = w.binItems[a][k];
					if (!c) continue;
					if (k <= b) { il += c, cl += w.binWeight[a][k]; for (int x = 0; x < 3; x++) ll[x] = fminf( ll[x], orderedFloat( w.binLo[a][k][x] ) ), lh[x] = fmaxf( lh[x], orderedFloat( w.binHi[a][k][x] ) ); }
					else { ir += c, cr += w.binWeight[a][k]; for (int x = 0; x < 3; x++) rl[x] = fminf( rl[x], orderedFloat( w.binLo[a][k][x] ) ), rh[x] = fmaxf( rh[x], orderedFloat( w.binHi[a][k][x] ) ); }
				}
				if (!il || !ir) continue;
				const float lx = lh[0] - ll[0], ly = lh[1] - ll[1], lz = lh[2] - ll[2], rx = rh[0] - rl[0], ry = rh[1] - rl[1], rz = rh[2] - rl[2];
				const float cost = (lx * ly + ly * lz + lz * lx) * (float)cl + (rx * ry + ry * rz + rz * rx) * (float)cr;
				if (cost < bestCost) bestCost = cost, bestCand = t, bestLeft = (uint32_t)il, bestLeftWeight = (uint32_t)cl;
			}
			for (int d = 16; d; d >>= 1)
			{
				const float oc = __shfl_xor_sync( 0xffffffffu, bestCost, d );
				const uint32_t ot = __shfl_xor_sync( 0xffffffffu, bestCand, d ), ol = __shfl_xor_sync( 0xffffffffu, bestLeft, d ), ow = __shfl_xor_sync( 0xffffffffu, bestLeftWeight, d );
				if (oc < bestCost || (oc == bestCost && ot < bestCand)) bestCost = oc, bestCand = ot, bestLeft = ol, bestLeftWeight = ow;
			}
			if (bestCand != 0xffffffffu)
			{
				// stable partition through the scratch copy
				const int a = bestCand / (SAHB_BINS - 1), b = bestCand % (SAHB_BINS - 1);
				nl = bestLeft, wl = bestLeftWeight;
				uint32_t doneL = 0, doneR = 0;
				for (uint32_t j0 = begin; j0 < begin + count; j0 += 32)
				{
					const uint32_t j = j0 + lane;
					const bool have = j < begin + count;
					const uint32_t o = have ? w.order[0][j] : 0;
					const bool left = have && max( 0, min( SAHB_BINS - 1, (int)((w.lo[a][o] + w.hi[a][o] - cmin[a]) * scale[a]) ) ) <= b;
					const uint32_t ml = __ballot_sync( 0xffffffffu, left ), mr = __ballot_sync( 0xffffffffu, have && !left ), below = (1u << lane) - 1u;
					if (have)
					{
						const uint32_t to = left ? begin + doneL + __popc( ml & below ) : begin + nl + doneR + __popc( mr & below );
						w.order[1][to] = (uint16_t)o;
					}
					doneL += __popc( ml ), doneR += __popc( mr );
				}
				__syncwarp();
				for (uint32_t j = begin + lane; j < begin + count; j += 32) w.order[0][j] = w.order[1][j];
				__syncwarp();
			}
		}
		if (wl == 0xffffffffu)
		{
			wl = 0;
			for (uint32_t j = begin + lane; j < begin + nl; j += 32) wl += w.weight[w.order[0][j]];
			for (int d = 16; d; d >>= 1) wl += __shfl_xor_sync( 0xffffffffu, wl, d );
		}
		// ---- the node's links: a side with one item is that item, the others take the next spare indices
		const uint32_t nr = count - nl, wr = weight - wl;
		uint32_t idL = 0, idR = 0;
		if (nl > 1) { while (w.pool[nextPool] == root || w.pool[nextPool] == spareEnd) nextPool++; idL = w.pool[nextPool++]; }
		if (nr > 1) { while (w.pool[nextPool] == root || w.pool[nextPool] == spareEnd) nextPool++; idR = w.pool[nextPool++]; }
		if (lane == 0)
		{
			const uint32_t L = nl > 1 ? idL : w.ref[w.order[0][begin]], R = nr > 1 ? idR : w.ref[w.order[0][begin + nl]];
			childL[id] = L, childR[id] = R, rangeCount[id] = weight;
			if (L & LEAF) parentLeaf[L & ~LEAF] = id; else parentInner[L] = id;
			if (R & LEAF) parentLeaf[R & ~LEAF] = id; else parentInner[R] = id;
		}
		// the smaller side first, the larger one waits: the stack stays below log2( SAHB_CLUSTER ) + 1 entries
		const bool leftFirst = nl <= nr;
		const uint32_t aId = leftFirst ? idL : idR, aBegin = leftFirst ? begin : begin + nl, aCount = leftFirst ? nl : nr, aWeight = leftFirst ? wl : wr;
		const uint32_t bId = leftFirst ? idR : idL, bBegin = leftFirst ? begin + nl : begin, bCount = leftFirst ? nr : nl, bWeight = leftFirst ? wr : wl;
		if (bCount > 1 && aCount > 1)
		{
			if (lane == 0) w.stackId[sp] = bId, w.stackBegin[sp] = bBegin, w.stackCount[sp] = bCount, w.stackWeight[sp] = bWeight;
			sp++;
			__syncwarp();
		}
		if (aCount > 1) id = aId, begin = aBegin, count = aCount, weight = aWeight;
		else if (bCount > 1) id = bId, begin = bBegin, count = bCount, weight = bWeight;
		else
		{
			if (!sp) break;
			sp--;
			id = w.stackId[sp], begin = w.stackBegin[sp], count = w.stackCount[sp], weight = w.stackWeight[sp];
		}
	}
}

// ---- 3d. Binned SAH from the top (builder 4): no space-filling curve at all.  Nodes with more than SAHB_CLUSTER triangles are split
// level by level -- every triangle of such a node adds its box to one of TOP_BINS bins per axis of its node (ordered-int atomics in
// global memory), a warp per node picks the split with the smallest sum of area * triangles, one running sum over all positions gives
// every triangle its place in a stable partition (left part first), and the children's centroid bounds are gathered while the triangles
// move.  Children with at most SAHB_CLUSTER triangles go to the list that sahbBuildKernel<false> finishes, one warp each, in shared
// memory.  This is the host builder's algorithm (bvh_builder.cpp: binned SAH over centroids, 64 bins on large nodes) in the shape the
// device wants; the trees are not the host's bit for bit (bin counts, tie order), but of the same quality.
#ifndef TOP_BINS_N
#define TOP_BINS_N 64
#endif
constexpr int TOP_BINS = TOP_BINS_N;
constexpr uint32_t TOP_NONE = 0xffffffffu;
struct TopNodes { uint32_t *id, *begin, *count; int* cb; };   // the large nodes of a level; cb = centroid bounds, six ordered ints each
struct TopBins { int *lo, *hi, *count; };                       // per node and axis and bin: box (3 + 3 ordered ints), triangles

__device__ __forceinline__ int topBinOf( const float c, const float cmin, const float cmax )
{
	const float scale = cmax > cmin ? (float)TOP_BINS * 0.999f / (cmax - cmin) : 0.f;
	return max( 0, min( TOP_BINS - 1, (int)((c - cmin) * scale) ) ); // (max: a centroid that is not a number must not name a bin outside the node's)
}

__global__ void topInitKernel( const uint32_t T, uint32_t* __restrict__ order, uint32_t* __restrict__ nodeOf, const TopNodes act, const int* __restrict__ centroidBounds, uint32_t* parentInner )
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < T) order[i] = i, nodeOf[i] = 0;
	if (i == 0)
	{
		act.id[0] = 0, act.begin[0] = 0, act.count[0] = T, parentInner[0] = 0xffffffffu;
		for (int k = 0; k < 6; k++) act.cb[k] = centroidBounds[k];
	}
}

__global__ void topClearBinsKernel( const uint32_t nAct, const TopBins bins, int* nextCb, const uint32_t nextCap )
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < nAct * 3 * TOP_BINS)
	{
		bins.count[i] = 0;
		for (int k = 0; k < 3; k++) bins.lo[i * 3 + k] = 0x7fffffff, bins.hi[i * 3 + k] = (int)0x80000000;
	}
	if (i < nextCap * 6) nextCb[i] = (i % 6) < 3 ? 0x7fffffff : (int)0x80000000;
}

// (Triangles that are neighbours in the mesh are neighbours here for the first levels: without care a warp sends 32 atomics to one
// address, a level of the sphere field took 8 ms.  So: the lanes of a warp that meet in a bin reduce among themselves first
// (__match_any_sync + __reduce_*_sync) and one of them speaks for all; and what goes to the node of the block's first triangle is
// collected in shared memory and flushed once per block.)
__global__ void __launch_bounds__( 256 ) topBinKernel( const uint32_t T, const uint32_t* __restrict__ order, const uint32_t* __restrict__ nodeOf, const TopNodes act,
	const float4* __restrict__ primLo, const float4* __restrict__ primHi, const TopBins bins )
{
	__shared__ int sLo[3 * TOP_BINS * 3], sHi[3 * TOP_BINS * 3], sCount[3 * TOP_BINS];
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	const uint32_t blockNode = nodeOf[blockIdx.x * blockDim.x];
	for (uint32_t t = threadIdx.x; t < 3 * TOP_BINS; t += blockDim.x)
	{
		sCount[t] = 0;
		for (int k = 0; k < 3; k++) sLo[t * 3 + k] = 0x7fffffff, sHi[t * 3 + k] = (int)0x80000000;
	}
	__syncthreads();
	const uint32_t a = i < T ? nodeOf[i] : TOP_NONE;
	const bool valid = a != TOP_NONE;
	float c[3] = { 0, 0, 0 };
	int ol[3] = { 0, 0, 0 }, oh[3] = { 0, 0, 0 };
	if (valid)
	{
		const uint32_t p = order[i];
		const float4 l = primLo[p], h = primHi[p];
		c[0] = 0.5f * (l.x + h.x), c[1] = 0.5f * (l.y + h.y), c[2] = 0.5f * (l.z + h.z);
		ol[0] = orderedInt( l.x ), ol[1] = orderedInt( l.y ), ol[2] = orderedInt( l.z ), oh[0] = orderedInt( h.x ), oh[1] = orderedInt( h.y ), oh[2] = orderedInt( h.z );
	}
	for (int ax = 0; ax < 3; ax++)
	{
		float cmin = 0, cmax = 0;
		if (valid) cmin = orderedFloat( act.cb[a * 6 + ax] ), cmax = orderedFloat( act.cb[a * 6 + 3 + ax] );
		const bool part = valid && cmax > cmin;
		const uint32_t pm = __ballot_sync( 0xffffffffu, part );
		if (!part) continue;
		const uint32_t b = (a * 3 + ax) * TOP_BINS + topBinOf( c[ax], cmin, cmax );
		const uint32_t group = __match_any_sync( pm, b );
		int gl[3], gh[3];
		for (int k = 0; k < 3; k++) gl[k] = __reduce_min_sync( group, ol[k] ), gh[k] = __reduce_max_sync( group, oh[k] );
		if ((threadIdx.x & 31u) != (uint32_t)__ffs( group ) - 1u) continue;
		if (a == blockNode)
		{
			const uint32_t sb = b - a * 3 * TOP_BINS;
			for (int k = 0; k < 3; k++) atomicMin( sLo + sb * 3 + k, gl[k] ), atomicMax( sHi + sb * 3 + k, gh[k] );
			atomicAdd( sCount + sb, __popc( group ) );
		}
		else
		{
			for (int k = 0; k < 3; k++) atomicMin( bins.lo + b * 3 + k, gl[k] ), atomicMax( bins.hi + b * 3 + k, gh[k] );
			atomicAdd( bins.count + b, __popc( group ) );
		}
	}
	__syncthreads();
	if (blockNode == TOP_NONE) return;
	for (uint32_t t = threadIdx.x; t < 3 * TOP_BINS; t += blockDim.x) if (sCount[t])
	{
		const uint32_t b = blockNode * 3 * TOP_BINS + t;
		for (int k = 0; k < 3; k++) atomicMin( bins.lo + b * 3 + k, sLo[t * 3 + k] ), atomicMax( bins.hi + b * 3 + k, sHi[t * 3 + k] );
		atomicAdd( bins.count + b, sCount[t] );
	}
}

// one warp per large node: the split, the children's indices, and where the children go (next level, the list of small subtrees, or a leaf)
__global__ void topSplitKernel( const uint32_t nAct, const TopNodes act, const TopBins bins, uint4* __restrict__ split, const TopNodes next, uint32_t* nextCount,
	uint4* __restrict__ smallRoots, uint32_t* smallCount, uint32_t* nodeCounter, uint32_t* __restrict__ childL, uint32_t* __restrict__ childR, uint32_t* __restrict__ rangeCount,
	uint32_t* __restrict__ parentInner, uint32_t* __restrict__ parentLeaf )
{
	const uint32_t a = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
	if (a >= nAct) return;
	const uint32_t id = act.id[a], begin = act.begin[a], count = act.count[a];
	float bestCost = 3.0e38f;
	uint32_t bestCand = 0xffffffffu, bestLeft = 0;
	// every lane holds TOP_BINS / 32 neighbouring bins of the axis; the union of the bins up to a candidate (from the left) and of the
	// bins behind it (from the right) are two warp scans, not a loop over the bins per candidate (which made this kernel the second longest
	// of the build while the first levels have 1, 2, 4 ... nodes: profiles/r02_ncu_build.txt).  min / max: the order of the unions does
	// not change a bit of them.
	constexpr int PER = TOP_BINS / 32;
	static_assert( TOP_BINS % 32 == 0, "topSplitKernel: a whole number of bins per lane" );
	struct Part { float lo[3], hi[3]; int n; };
	auto join = []( Part& x, const Part& y ) { for (int k = 0; k < 3; k++) x.lo[k] = fminf( x.lo[k], y.lo[k] ), x.hi[k] = fmaxf( x.hi[k], y.hi[k] ); x.n += y.n; };
	auto shift = [lane]( const Part& x, const int d, const bool up ) {
		Part y;
		for (int k = 0; k < 3; k++)
			y.lo[k] = up ? __shfl_up_sync( 0xffffffffu, x.lo[k], d ) : __shfl_down_sync( 0xffffffffu, x.lo[k], d ),
			y.hi[k] = up ? __shfl_up_sync( 0xffffffffu, x.hi[k], d ) : __shfl_down_sync( 0xffffffffu, x.hi[k], d );
		y.n = up ? __shfl_up_sync( 0xffffffffu, x.n, d ) : __shfl_down_sync( 0xffffffffu, x.n, d );
		if (up ? (int)lane < d : (int)lane + d > 31) { for (int k = 0; k < 3; k++) y.lo[k] = 3.0e38f, y.hi[k] = -3.0e38f; y.n = 0; }
		return y;
	};
	for (uint32_t ax = 0; ax < 3; ax++)
	{
		if (!(orderedFloat( act.cb[a * 6 + 3 + ax] ) > orderedFloat( act.cb[a * 6 + ax] ))) continue;
		const uint32_t base = (a * 3 + ax) * TOP_BINS + lane * PER;
		Part bin[PER], left[PER], right[PER];
		for (int k = 0; k < PER; k++)
		{
			bin[k].n = bins.count[base + k];
			for (int x = 0; x < 3; x++)
				bin[k].lo[x] = bin[k].n ? orderedFloat( bins.lo[(base + k) * 3 + x] ) : 3.0e38f, bin[k].hi[x] = bin[k].n ? orderedFloat( bins.hi[(base + k) * 3 + x] ) : -3.0e38f;
		}
		// left[k] = bins 0 .. this lane's k-th, right[k] = this lane's k-th .. the last bin
		left[0] = bin[0];
		for (int k = 1; k < PER; k++) left[k] = left[k - 1], join( left[k], bin[k] );
		right[PER - 1] = bin[PER - 1];
		for (int k = PER - 2; k >= 0; k--) right[k] = right[k + 1], join( right[k], bin[k] );
		Part fromLeft = left[PER - 1], fromRight = right[0]; // inclusive scans over the lanes' totals ...
		for (int d = 1; d < 32; d <<= 1) { const Part l = shift( fromLeft, d, true ), r = shift( fromRight, d, false ); join( fromLeft, l ), join( fromRight, r ); }
		const Part before = shift( fromLeft, 1, true ), after = shift( fromRight, 1, false ); // ... made exclusive
		for (int k = 0; k < PER; k++) join( left[k], before ), join( right[k], after );
		const Part nextRight = shift( right[0], 1, false ); // the suffix that starts at the next lane's first bin
		for (int k = 0; k < PER; k++)
		{
			const uint32_t g = lane * PER + k; // the bins up to and including g go left
			if (g >= TOP_BINS - 1) continue;
			const Part& L = left[k];
			const Part& R = k + 1 < PER ? right[k + 1 < PER ? k + 1 : 0] : nextRight;
			if (!L.n || !R.n) continue;
			const float lx = L.hi[0] - L.lo[0], ly = L.hi[1] - L.lo[1], lz = L.hi[2] - L.lo[2], rx = R.hi[0] - R.lo[0], ry = R.hi[1] - R.lo[1], rz = R.hi[2] - R.lo[2];
			const float cost = (lx * ly + ly * lz + lz * lx) * (float)L.n + (rx * ry + ry * rz + rz * rx) * (float)R.n;
			const uint32_t t = ax * (TOP_BINS - 1) + g;
			if (cost < bestCost || (cost == bestCost && t < bestCand)) bestCost = cost, bestCand = t, bestLeft = (uint32_t)L.n;
		}
	}
	for (int d = 16; d; d >>= 1)
	{
		const float oc = __shfl_xor_sync( 0xffffffffu, bestCost, d );
		const uint32_t ot = __shfl_xor_sync( 0xffffffffu, bestCand, d ), ol = __shfl_xor_sync( 0xffffffffu, bestLeft, d );
		if (oc < bestCost || (oc == bestCost && ot < bestCand)) bestCost = oc, bestCand = ot, bestLeft = ol;
	}
	if (lane) return;
	const uint32_t nl = bestCand != 0xffffffffu ? bestLeft : count >> 1, nr = count - nl; // (no split separates the centroids: halves in the order they are in)
	uint32_t ref[2], where[2];
	for (int k = 0; k < 2; k++)
	{
		const uint32_t n = k ? nr : nl, b = k ? begin + nl : begin;
		where[k] = TOP_NONE;
		if (n == 1) { ref[k] = LEAF | b, parentLeaf[b] = id; continue; }
		const uint32_t c = atomicAdd( nodeCounter, 1u );
		ref[k] = c, parentInner[c] = id;
		if (n <= SAHB_CLUSTER) smallRoots[atomicAdd( smallCount, 1u )] = make_uint4( c, b, n, 0 );
		else
		{
			const uint32_t x = atomicAdd( nextCount, 1u );
			next.id[x] = c, next.begin[x] = b, next.count[x] = n, where[k] = x;
		}
	}
	childL[id] = ref[0], childR[id] = ref[1], rangeCount[id] = count;
	split[a] = make_uint4( bestCand, nl, where[0], where[1] );
}

__device__ __forceinline__ bool topGoesLeft( const uint32_t i, const uint32_t a, const uint32_t* __restrict__ order, const TopNodes& act, const uint4 sp,
	const float4* __restrict__ primLo, const float4* __restrict__ primHi, float c[3] )
{
	const uint32_t p = order[i];
	const float4 l = primLo[p], h = primHi[p];
	c[0] = 0.5f * (l.x + h.x), c[1] = 0.5f * (l.y + h.y), c[2] = 0.5f * (l.z + h.z);
	if (sp.x == 0xffffffffu) return i - act.begin[a] < sp.y;
	const uint32_t ax = sp.x / (TOP_BINS - 1), b = sp.x % (TOP_BINS - 1);
	return (uint32_t)topBinOf( c[ax], orderedFloat( act.cb[a * 6 + ax] ), orderedFloat( act.cb[a * 6 + 3 + ax] ) ) <= b;
}

__global__ void topFlagKernel( const uint32_t T, const uint32_t* __restrict__ order, const uint32_t* __restrict__ nodeOf, const TopNodes act, const uint4* __restrict__ split,
	const float4* __restrict__ primLo, const float4* __restrict__ primHi, uint32_t* __restrict__ flag )
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= T) return;
	const uint32_t a = nodeOf[i];
	float c[3];
	flag[i] = a != TOP_NONE && topGoesLeft( i, a, order, act, split[a], primLo, primHi, c ) ? 1u : 0u;
}

__global__ void topScatterKernel( const uint32_t T, const uint32_t* __restrict__ order, const uint32_t* __restrict__ nodeOf, const TopNodes act, const uint4* __restrict__ split,
	const float4* __restrict__ primLo, const float4* __restrict__ primHi, const uint32_t* __restrict__ leftBefore, uint32_t* __restrict__ orderOut, uint32_t* __restrict__ nodeOfOut, int* nextCb )
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= T) return;
	const uint32_t a = nodeOf[i];
	if (a == TOP_NONE) { orderOut[i] = order[i], nodeOfOut[i] = TOP_NONE; return; }
	const uint4 sp = split[a];
	float c[3];
	const bool left = topGoesLeft( i, a, order, act, sp, primLo, primHi, c );
	const uint32_t begin = act.begin[a], rank = leftBefore[i] - leftBefore[begin];
	const uint32_t to = left ? begin + rank : begin + sp.y + (i - begin - rank), where = left ? sp.z : sp.w;
	orderOut[to] = order[i], nodeOfOut[to] = where;
	// (the lanes of a warp that go to the same child speak through one of them, as in topBinKernel)
	const uint32_t group = __match_any_sync( __activemask(), where );
	int gl[3], gh[3];
	for (int k = 0; k < 3; k++) gl[k] = __reduce_min_sync( group, orderedInt( c[k] ) ), gh[k] = __reduce_max_sync( group, orderedInt( c[k] ) );
	if (where != TOP_NONE && (threadIdx.x & 31u) == (uint32_t)__ffs( group ) - 1u)
		for (int k = 0; k < 3; k++) atomicMin( nextCb + where * 6 + k, gl[k] ), atomicMax( nextCb + where * 6 + 3 + k, gh[k] );
}

// ---- 3b / 4b. agglomerative alternative to the radix tree: PLOC (Meister, Bittner: "Parallel Locally-Ordered Clustering for Bounding Volume
// Hierarchy Construction", TVCG 2018).  The clusters -- at first the triangles in Morton order -- each look for the neighbour within
// `radius` places along the curve whose union with them has the smallest surface; pairs that chose each other merge into a binary node;
// the survivors are compacted, still in curve order, and the round repeats until one cluster is left.  The tree is built bottom-up by
// surface area instead of top-down by key bits: the Morton order only decides who may meet whom.  Binary nodes are numbered from the
// back (the last merge, the root, gets index 0, where the collapse starts) and receive box, primitive count and collapse plan as they
// are made -- the jobs of radixTreeKernel and binaryBoxesKernel.
struct PlocClusters { uint32_t* ref; float4 *lo, *hi; }; // ref: LEAF | sorted position, or a binary node

__global__ void plocInitKernel( const uint32_t* __restrict__ sortedPrim, const float4* __restrict__ primLo, const float4* __restrict__ primHi, const uint32_t n, const PlocClusters c )
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const uint32_t p = sortedPrim[i];
	c.ref[i] = LEAF | i, c.lo[i] = primLo[p], c.hi[i] = primHi[p];
}

template <int RADIUS> __global__ void __launch_bounds__( 256 ) plocNearestKernel( const PlocClusters c, const uint32_t* __restrict__ nPtr, uint32_t* __restrict__ nearest )
{
	__shared__ float4 sLo[256 + 2 * RADIUS], sHi[256 + 2 * RADIUS];
	const uint32_t n = *nPtr;
	const int blockFirst = (int)(blockIdx.x * 256);
	if (blockFirst >= (int)n) return;
	for (int t = threadIdx.x; t < 256 + 2 * RADIUS; t += 256)
	{
		const int j = blockFirst - RADIUS + t;
		if (j >= 0 && j < (int)n) sLo[t] = c.lo[j], sHi[t] = c.hi[j];
	}
	__syncthreads();
	const int i = blockFirst + (int)threadIdx.x;
	if (i >= (int)n) return;
	const float4 lo = sLo[threadIdx.x + RADIUS], hi = sHi[threadIdx.x + RADIUS];
	// Among equal areas the nearer place wins, and among the two places at distance d the one that pairs the clusters up in aligned
	// blocks of d (even block: the partner above, odd block: below).  That order guarantees that the smallest pair of a round is mutual
	// (the build always advances) and that a run of identical boxes -- coincident triangles -- halves per round instead of losing one
	// pair per round.
	float best = 3.0e38f;
	int bestJ = -1;
#pragma unroll
	for (int d = 1; d <= RADIUS; d++)
	{
		const int up = ((i / d) & 1) == 0 ? d : -d;
#pragma unroll
		for (int side = 0; side < 2; side++)
		{
			const int o = side ? -up : up, j = i + o;
			if (j < 0 || j >= (int)n) continue;
			const float4 l = sLo[(int)threadIdx.x + RADIUS + o], h = sHi[(int)threadIdx.x + RADIUS + o];
			const float x = fmaxf( hi.x, h.x ) - fminf( lo.x, l.x ), y = fmaxf( hi.y, h.y ) - fminf( lo.y, l.y ), z = fmaxf( hi.z, h.z ) - fminf( lo.z, l.z );
			const float area = x * y + y * z + z * x;
			if (area < best) best = area, bestJ = j;
		}
	}
	nearest[i] = (uint32_t)bestJ;
}

// pairs that chose each other merge (the lower index does the work and keeps the place); flag = the cluster lives on
__global__ void plocMergeKernel( const PlocClusters c, const uint32_t* __restrict__ nPtr, const uint32_t* __restrict__ nearest, uint32_t* __restrict__ flag,
	uint32_t* __restrict__ mergeCounter, const uint32_t T, const uint32_t* __restrict__ sortedPrim, const float4* __restrict__ primLo, const float4* __restrict__ primHi,
	uint32_t* __restrict__ childL, uint32_t* __restrict__ childR, uint32_t* __restrict__ rangeCount, float4* __restrict__ nodeLo, float4* __restrict__ nodeHi,
	float4* __restrict__ planLo, float4* __restrict__ planHi, const float primCost )
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x, n = *nPtr;
	if (i >= n) return;
	const uint32_t j = nearest[i];
	if (j >= n || nearest[j] != i) { flag[i] = 1; return; }
	if (j < i) { flag[i] = 0; return; }
	const uint32_t node = T - 2 - atomicAdd( mergeCounter, 1u );
	const uint32_t L = c.ref[i], R = c.ref[j];
	const float4 ll = c.lo[i], lh = c.hi[i], rl = c.lo[j], rh = c.hi[j];
	const float4 nl = make_float4( fminf( ll.x, rl.x ), fminf( ll.y, rl.y ), fminf( ll.z, rl.z ), 0 ), nh = make_float4( fmaxf( lh.x, rh.x ), fmaxf( lh.y, rh.y ), fmaxf( lh.z, rh.z ), 0 );
	const uint32_t count = ((L & LEAF) ? 1u : rangeCount[L]) + ((R & LEAF) ? 1u : rangeCount[R]);
	childL[node] = L, childR[node] = R, rangeCount[node] = count, nodeLo[node] = nl, nodeHi[node] = nh;
	float cl[7], cr[7], cost[7];
	loadPlan( L, sortedPrim, primLo, primHi, planLo, planHi, primCost, cl );
	loadPlan( R, sortedPrim, primLo, primHi, planLo, planHi, primCost, cr );
	const float ex = nh.x - nl.x, ey = nh.y - nl.y, ez = nh.z - nl.z;
	uint32_t bits;
	mergePlan( cl, cr, ex * ey + ey * ez + ez * ex, count, 3, primCost, cost, bits );
	planLo[node] = make_float4( cost[0], cost[1], cost[2], cost[3] ), planHi[node] = make_float4( cost[4], cost[5], cost[6], __uint_as_float( bits ) );
	c.ref[i] = node, c.lo[i] = nl, c.hi[i] = nh, flag[i] = 1;
}

// survivors move up, in order (offsets = exclusive sum of the flags); the last thread leaves the new count
__global__ void plocCompactKernel( const PlocClusters in, const PlocClusters out, uint32_t* __restrict__ nPtr, const uint32_t nOld, const uint32_t* __restrict__ flag,
	const uint32_t* __restrict__ offset )
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= nOld) return;
	if (flag[i]) { const uint32_t o = offset[i]; out.ref[o] = in.ref[i], out.lo[o] = in.lo[i], out.hi[o] = in.hi[i]; }
	if (i == nOld - 1) *nPtr = offset[i] + flag[i];
}

// the (at most three) primitives of a subtree that became a leaf slot, left to right.  (For the radix tree that is a contiguous run of the
// sorted order; the agglomerative builder below merges clusters that are near each other but not adjacent, so the tree is walked.)
__device__ __forceinline__ uint32_t leafSlotPrims( const uint32_t* __restrict__ childL, const uint32_t* __restrict__ childR, const uint32_t ref, uint32_t out[3] )
{
	if (ref & LEAF) { out[0] = ref & ~LEAF; return 1; }
	uint32_t n = 0;
	const uint32_t c[2] = { childL[ref], childR[ref] };
#pragma unroll
	for (int k = 0; k < 2; k++)
	{
		if (c[k] & LEAF) { if (n < 3) out[n] = c[k] & ~LEAF; n++; }
		else
		{
			const uint32_t l = childL[c[k]], r = childR[c[k]]; // a subtree of two: both are leaves
			if (n < 3) out[n] = l & ~LEAF;
			n++;
			if (n < 3) out[n] = r & ~LEAF;
			n++;
		}
	}
	return n < 3 ? n : 3;
}

struct CollapseArgs {
	const uint32_t *sortedPrim, *childL, *childR, *rangeCount;
	const float4 *primLo, *primHi, *nodeLo, *nodeHi, *planHi;
	uint32_t* wideToBinary;   // [maxNodes] binary inner node that wide node w represents
	uint4* nodes;             // output, relative index 0 = this tree's root (pool pointer already offset)
	uint32_t* primOut;        // output: original primitive index of leaf-order slot j goes to primOut[j * primOutStride] (triangle records: the .w of the
	uint32_t primOutStride;   //   first float4, stride 12 words; the top level: a plain array of instance ids)
	uint32_t* counters;       // [0] wide nodes allocated, [1] leaf-order triangle slots allocated, [2] overflow flag
	uint32_t nodeBase, triBase, maxNodes;
	// deterministic placement (the mesh builder): levelTotals[i] = (inner children << 32 | leaf primitives) of all wide nodes of the level before the i-th;
	// null = children and primitives are claimed with atomicAdd in whatever order the groups arrive (the one-block top-level build)
	const unsigned long long* levelTotals;
	uint32_t levelBegin, levelChildBase, levelPrimBase;
};

__device__ __forceinline__ void refBox( const CollapseArgs& a, const uint32_t ref, float lo[3], float hi[3], uint32_t& count )
{
	float4 l, h;
	if (ref & LEAF) { const uint32_t p = a.sortedPrim[ref & ~LEAF]; l = a.primLo[p], h = a.primHi[p], count = 1; }
	else l = a.nodeLo[ref], h = a.nodeHi[ref], count = a.rangeCount[ref];
	lo[0] = l.x, lo[1] = l.y, lo[2] = l.z, hi[0] = h.x, hi[1] = h.y, hi[2] = h.z;
}

// 5. one level of the top-down collapse: thread w builds wide node `levelBegin + w` from its binary node
#if !B200_USE_GROUP_COLLAPSE
__device__ void collapseNode( const CollapseArgs& a, const uint32_t w )
{
	const uint32_t root = a.wideToBinary[w];
	uint32_t ref[8], cnt[8];
	float lo[8][3], hi[8][3];
	bool innerChild[8];
	int n = 0;
#if B200_DP_COLLAPSE
	// the children are the forest the plan chose: left gets k8 of the eight slots, right the rest; a subtree with an allowance of i
	// roots either settles for fewer (choice 0), becomes one root (a leaf slot or a wide node), or hands k roots to its left child
	{
		const uint32_t rootBits = __float_as_uint( a.planHi[root].w ), k8 = (rootBits >> 18) & 7u;
		uint32_t stackRef[8], stackAllow[8];
		int sp = 0;
		stackRef[sp] = a.childR[root], stackAllow[sp++] = 8 - k8;
		stackRef[sp] = a.childL[root], stackAllow[sp++] = k8;
		while (sp)
		{
			const uint32_t m = stackRef[--sp];
			uint32_t allow = stackAllow[sp];
			uint32_t bits = 0;
			if (!(m & LEAF))
			{
				bits = __float_as_uint( a.planHi[m].w );
				while (allow > 1 && ((bits >> (3 * (allow - 2))) & 7u) == 0) allow--;
			}
			if ((m & LEAF) || allow == 1)
			{
				ref[n] = m, innerChild[n] = !(m & LEAF) && !((bits >> 21) & 1u);
				refBox( a, m, lo[n], hi[n], cnt[n] );
				n++;
			}
			else
			{
				const uint32_t k = (bits >> (3 * (allow - 2))) & 7u;
				stackRef[sp] = a.childR[m], stackAllow[sp++] = allow - k;
				stackRef[sp] = a.childL[m], stackAllow[sp++] = k;
			}
		}
	}
#else
	n = 2;
	ref[0] = a.childL[root], ref[1] = a.childR[root];
	refBox( a, ref[0], lo[0], hi[0], cnt[0] ), refBox( a, ref[1], lo[1], hi[1], cnt[1] );
	// open the largest child that is still too big for a leaf slot until all eight slots are used
	while (n < 8)
	{
		int best = -1;
		float bestArea = -1.f;
		for (int i = 0; i < n; i++) if (cnt[i] > 3)
		{
			const float x = hi[i][0] - lo[i][0], y = hi[i][1] - lo[i][1], z = hi[i][2] - lo[i][2], area = x * y + y * z + z * x;
			if (area > bestArea) bestArea = area, best = i;
		}
		if (best < 0) break;
		const uint32_t opened = ref[best];
		ref[best] = a.childL[opened], ref[n] = a.childR[opened];
		refBox( a, ref[best], lo[best], hi[best], cnt[best] ), refBox( a, ref[n], lo[n], hi[n], cnt[n] );
		n++;
	}
	for (int i = 0; i < n; i++) innerChild[i] = cnt[i] > 3;
#endif
	// octant slots: slot s is visited first by rays whose direction signs are the complement of s (traverse.cu), so a child
	// goes to the slot whose sign vector it lies farthest against -- greedy over the 8 x n table, as AssignSlots does on the host
	float nlo[3], nhi[3];
	for (int c = 0; c < 3; c++)
	{
		nlo[c] = lo[0][c], nhi[c] = hi[0][c];
		for (int i = 1; i < n; i++) nlo[c] = fminf( nlo[c], lo[i][c] ), nhi[c] = fmaxf( nhi[c], hi[i][c] );
	}
	// (the child's offset from the node centre is computed once per child and the eight signed sums once per (slot, child) pair; the n
	// greedy rounds then only compare.  Recomputing them inside the rounds -- 8 x n x n x 3 steps -- made this loop a quarter of all
	// instructions of the one-block top-level build, where the first levels of the tree have 1, 8, ... nodes and as many busy threads.
	// Same sums, same scan order, same strict <: the slots, and with them the trees, are the ones the first version chose.)
	int childInSlot[8] = { -1, -1, -1, -1, -1, -1, -1, -1 };
	float score[8][8];
	for (int i = 0; i < n; i++)
	{
		float rel[3];
		for (int c = 0; c < 3; c++) rel[c] = 0.5f * (lo[i][c] + hi[i][c]) - 0.5f * (nlo[c] + nhi[c]);
		for (int s = 0; s < 8; s++)
		{
			float d = 0;
			for (int c = 0; c < 3; c++) d += ((s >> (2 - c)) & 1) ? -rel[c] : rel[c];
			score[s][i] = d;
		}
	}
	uint32_t done = 0, freeSlots = 0xffu;
	for (int k = 0; k < n; k++)
	{
		float best = 3.0e38f;
		int bs = 0, bi = 0;
		for (int s = 0; s < 8; s++) if (freeSlots & (1u << s))
			for (int i = 0; i < n; i++) if (!(done & (1u << i)) && score[s][i] < best) best = score[s][i], bs = s, bi = i;
		childInSlot[bs] = bi, done |= 1u << bi, freeSlots &= ~(1u << bs);
	}
	uint32_t innerCount = 0, leafPrims = 0;
	for (int i = 0; i < n; i++) if (innerChild[i]) innerCount++; else leafPrims += cnt[i];
	const uint32_t childBase = innerCount ? atomicAdd( a.counters, innerCount ) : 0;
	const uint32_t primBase = leafPrims ? atomicAdd( a.counters + 1, leafPrims ) : 0;
	if (childBase + innerCount > a.maxNodes) { atomicExch( a.counters + 2, 1u ); return; }
	uint32_t imask = 0, metaLo = 0, metaHi = 0, inner = 0, off = 0;
	for (int s = 0; s < 8; s++)
	{
		const int i = childInSlot[s];
		if (i < 0) continue;
		uint32_t meta;
		if (innerChild[i])
		{
			imask |= 1u << s, meta = 0x20u | (24u + (uint32_t)s);
			a.wideToBinary[childBase + inner] = ref[i];
			inner++;
		}
		else
		{
			meta = (((1u << cnt[i]) - 1u) << 5) | off;
			// the subtree's triangles are a contiguous run of the sorted order
			uint32_t at[3];
			leafSlotPrims( a.childL, a.childR, ref[i], at );
			for (uint32_t k = 0; k < cnt[i]; k++)
				a.primOut[(size_t)(primBase + off + k) * a.primOutStride] = a.sortedPrim[at[k]];
			off += cnt[i];
		}
		if (s < 4) metaLo |= meta << (8 * s); else metaHi |= meta << (8 * (s - 4));
	}
	uint4* np = a.nodes + (size_t)w * 5;
	np[0] = make_uint4( 0, 0, 0, imask << 24 );
	np[1] = make_uint4( a.nodeBase + childBase, a.triBase + primBase, metaLo, metaHi );
	np[2] = np[3] = np[4] = make_uint4( 0, 0, 0, 0 );
}
#endif

// The same step by EIGHT LANES per wide node: lane i of the group holds child i in registers, what collapseNode keeps in dynamically
// indexed local arrays (lo / hi / score / stack: hundreds of dependent local-memory accesses per node, which is what the one-block
// top-level build spent two thirds of its time in -- its first levels have 1, 8, ... nodes and as many busy threads).  Same forest walk,
// same sums in the same order, same strict-< scan order of the greedy slot assignment: the nodes are bit for bit those of
// collapseNode (tests/test_parity_gpu.py holds the device-built top level against the host builder's).
// Called by all eight lanes of a group together; `gmask` names them in the warp, `sub` = lane within the group.
// the forest the plan chose for wide node w (see collapseNode); the walk is the same in all eight lanes of the group, stack entry j lives in lane j,
// child n goes to lane n.  Returns the number of children.
__device__ __forceinline__ uint32_t forestWalk( const CollapseArgs& a, const uint32_t w, const uint32_t sub, const uint32_t gmask, const uint32_t gbase, uint32_t& myRef, uint32_t& myBits )
{
	const uint32_t root = a.wideToBinary[w];
	uint32_t stRef = 0, stAllow = 0;
	uint32_t n = 0, sp = 2;
	{
		const uint32_t k8 = (__float_as_uint( a.planHi[root].w ) >> 18) & 7u;
		if (sub == 0) stRef = a.childR[root], stAllow = 8 - k8;
		if (sub == 1) stRef = a.childL[root], stAllow = k8;
	}
	while (sp)
	{
		--sp;
		const uint32_t m = __shfl_sync( gmask, stRef, gbase + sp );
		uint32_t allow = __shfl_sync( gmask, stAllow, gbase + sp ), bits = 0;
		if (!(m & LEAF))
		{
			bits = __float_as_uint( a.planHi[m].w );
			while (allow > 1 && ((bits >> (3 * (allow - 2))) & 7u) == 0) allow--;
		}
		if ((m & LEAF) || allow == 1)
		{
			if (sub == n) myRef = m, myBits = bits;
			n++;
		}
		else
		{
			const uint32_t k = (bits >> (3 * (allow - 2))) & 7u, right = a.childR[m], left = a.childL[m];
			if (sub == sp) stRef = right, stAllow = allow - k;
			sp++;
			if (sub == sp) stRef = left, stAllow = k;
			sp++;
		}
	}
	return n;
}

// first pass of a level in the deterministic mode: how many inner children and leaf primitives wide node w will claim
__device__ void collapseCountGroup( const CollapseArgs& a, const uint32_t w, const uint32_t sub, const uint32_t gmask, unsigned long long* __restrict__ out )
{
	const uint32_t gbase = (threadIdx.x & 31u) - sub;
	uint32_t myRef = 0, myBits = 0;
	const uint32_t n = forestWalk( a, w, sub, gmask, gbase, myRef, myBits );
	const bool have = sub < n, inner = have && !(myRef & LEAF) && !((myBits >> 21) & 1u);
	uint32_t leafPrims = (have && !inner) ? ((myRef & LEAF) ? 1u : a.rangeCount[myRef]) : 0u;
#pragma unroll
	for (int d = 1; d < 8; d <<= 1) leafPrims += __shfl_xor_sync( gmask, leafPrims, d );
	const uint32_t innerCount = __popc( (__ballot_sync( gmask, inner ) >> gbase) & 0xffu );
	if (sub == 0) out[w - a.levelBegin] = ((unsigned long long)innerCount << 32) | leafPrims;
}

__device__ void collapseNodeGroup( const CollapseArgs& a, const uint32_t w, const uint32_t sub, const uint32_t gmask )
{
	const uint32_t gbase = (threadIdx.x & 31u) - sub;
	uint32_t myRef = 0, myBits = 0;
	const uint32_t n = forestWalk( a, w, sub, gmask, gbase, myRef, myBits );
	const bool have = sub < n;
	const bool inner = have && !(myRef & LEAF) && !((myBits >> 21) & 1u);
	float lo[3] = { 3.0e38f, 3.0e38f, 3.0e38f }, hi[3] = { -3.0e38f, -3.0e38f, -3.0e38f };
	uint32_t cnt = 0;
	if (have) refBox( a, myRef, lo, hi, cnt );
	// node box: min / max over the children (any order gives the same floats)
	float nlo[3], nhi[3];
#pragma unroll
	for (int c = 0; c < 3; c++)
	{
		nlo[c] = lo[c], nhi[c] = hi[c];
#pragma unroll
		for (int d = 1; d < 8; d <<= 1) nlo[c] = fminf( nlo[c], __shfl_xor_sync( gmask, nlo[c], d ) ), nhi[c] = fmaxf( nhi[c], __shfl_xor_sync( gmask, nhi[c], d ) );
	}
	// my child's eight signed sums, then n greedy rounds: the smallest (score, slot, child) in the scan order of collapseNode
	float score[8];
	{
		float rel[3];
#pragma unroll
		for (int c = 0; c < 3; c++) rel[c] = 0.5f * (lo[c] + hi[c]) - 0.5f * (nlo[c] + nhi[c]);
#pragma unroll
		for (int s = 0; s < 8; s++)
		{
			float d = 0;
#pragma unroll
			for (int c = 0; c < 3; c++) d += ((s >> (2 - c)) & 1) ? -rel[c] : rel[c];
			score[s] = d;
		}
	}
	uint32_t freeSlots = 0xffu, mySlot = 8;
	bool done = !have;
	for (uint32_t k = 0; k < n; k++)
	{
		float best = 3.0e38f;
		uint32_t key = 0xffu; // slot << 3 | child: the scan order among equal scores
#pragma unroll
		for (int s = 0; s < 8; s++) if (!done && (freeSlots & (1u << s)) && score[s] < best) best = score[s], key = ((uint32_t)s << 3) | sub;
#pragma unroll
		for (int d = 1; d < 8; d <<= 1)
		{
			const float ob = __shfl_xor_sync( gmask, best, d );
			const uint32_t ok = __shfl_xor_sync( gmask, key, d );
			if (ob < best || (ob == best && ok < key)) best = ob, key = ok;
		}
		const uint32_t bs = (key >> 3) & 7u, bi = key & 7u;
		if (sub == bi) mySlot = bs, done = true;
		freeSlots &= ~(1u << bs);
	}
	// children and primitives of this node: one claim each for the group
	const uint32_t innerBits = (__ballot_sync( gmask, inner ) >> gbase) & 0xffu, innerCount = __popc( innerBits );
	uint32_t leafPrims = inner ? 0u : cnt;
#pragma unroll
	for (int d = 1; d < 8; d <<= 1) leafPrims += __shfl_xor_sync( gmask, leafPrims, d );
	uint32_t childBase = 0, primBase = 0;
	if (sub == 0)
	{
		if (a.levelTotals)
		{
			// the level's nodes in order: what the nodes before this one claim is known (collapseCountGroup + a running sum)
			const unsigned long long before = a.levelTotals[w - a.levelBegin];
			childBase = a.levelChildBase + (uint32_t)(before >> 32), primBase = a.levelPrimBase + (uint32_t)before;
			atomicMax( a.counters, childBase + innerCount ), atomicMax( a.counters + 1, primBase + leafPrims ); // (the level's last node leaves the totals)
		}
		else
		{
			childBase = innerCount ? atomicAdd( a.counters, innerCount ) : 0;
			primBase = leafPrims ? atomicAdd( a.counters + 1, leafPrims ) : 0;
		}
	}
	childBase = __shfl_sync( gmask, childBase, gbase ), primBase = __shfl_sync( gmask, primBase, gbase );
	if (childBase + innerCount > a.maxNodes) { if (sub == 0) atomicExch( a.counters + 2, 1u ); return; }
	// place in slot order: inner children before mine, leaf primitives before mine
	const uint32_t packed = have ? (0x80000000u | mySlot | (inner ? 8u : 0u) | (cnt << 4)) : 0u;
	uint32_t innerBefore = 0, off = 0;
#pragma unroll
	for (int j = 0; j < 8; j++)
	{
		const uint32_t o = __shfl_sync( gmask, packed, gbase + j );
		if ((o & 0x80000000u) && (o & 7u) < mySlot) { if (o & 8u) innerBefore++; else off += (o >> 4) & 0x07ffffffu; }
	}
	uint32_t imask = 0, metaLo = 0, metaHi = 0;
	if (have)
	{
		uint32_t meta;
		if (inner)
		{
			imask = 1u << mySlot, meta = 0x20u | (24u + mySlot);
			a.wideToBinary[childBase + innerBefore] = myRef;
		}
		else
		{
			meta = (((1u << cnt) - 1u) << 5) | off;
			uint32_t at[3];
			leafSlotPrims( a.childL, a.childR, myRef, at );
			for (uint32_t k = 0; k < cnt; k++) a.primOut[(size_t)(primBase + off + k) * a.primOutStride] = a.sortedPrim[at[k]];
		}
		if (mySlot < 4) metaLo = meta << (8 * mySlot); else metaHi = meta << (8 * (mySlot - 4));
	}
#pragma unroll
	for (int d = 1; d < 8; d <<= 1)
		imask |= __shfl_xor_sync( gmask, imask, d ), metaLo |= __shfl_xor_sync( gmask, metaLo, d ), metaHi |= __shfl_xor_sync( gmask, metaHi, d );
	if (sub == 0)
	{
		uint4* np = a.nodes + (size_t)w * 5;
		np[0] = make_uint4( 0, 0, 0, imask << 24 );
		np[1] = make_uint4( a.nodeBase + childBase, a.triBase + primBase, metaLo, metaHi );
		np[2] = np[3] = np[4] = make_uint4( 0, 0, 0, 0 );
	}
}

__global__ void collapseCountKernel( const CollapseArgs a, const uint32_t levelBegin, const uint32_t levelEnd, unsigned long long* __restrict__ out )
{
	const uint32_t w = levelBegin + (blockIdx.x * blockDim.x + threadIdx.x) / 8, sub = threadIdx.x & 7u;
	if (w < levelEnd) collapseCountGroup( a, w, sub, 0xffu << ((threadIdx.x & 31u) - sub), out );
}

__global__ void collapseKernel( const CollapseArgs a, const uint32_t levelBegin, const uint32_t levelEnd )
{
#if B200_USE_GROUP_COLLAPSE
	const uint32_t w = levelBegin + (blockIdx.x * blockDim.x + threadIdx.x) / 8, sub = threadIdx.x & 7u;
	if (w < levelEnd) collapseNodeGroup( a, w, sub, 0xffu << ((threadIdx.x & 31u) - sub) );
#else
	const uint32_t w = levelBegin + blockIdx.x * blockDim.x + threadIdx.x;
	if (w < levelEnd) collapseNode( a, w );
#endif
}


// ---- the top level, built by ONE thread block ------------------------------------------------------------------------------------
// Replaces the per-frame host rebuild (TopLevelBVH::Rebuild, lib/RenderCore_Bart/bvh.cpp:373, called from Render, rendercore.cpp:199;
// optixAccelBuild of the IAS, lib/rendercore_optix7/rendercore.cpp:426) for scenes of up to MAX_DEVICE_TLAS_INSTANCES instances: world
// boxes -> Morton keys -> bitonic sort -> Karras radix tree -> bottom-up boxes + collapse plan -> 8-wide collapse -> frames and
// quantised planes -> instance records in leaf order.  Everything a pass needs at once sits in shared memory; the passes are
// separated by block barriers, so the whole build is one launch of a few tens of microseconds and the host neither builds nor waits.
__device__ __forceinline__ uint32_t spread10( uint32_t x )
{
	x &= 0x3ffu;
	x = (x | x << 16) & 0x030000ffu;
	x = (x | x << 8) & 0x0300f00fu;
	x = (x | x << 4) & 0x030c30c3u;
	x = (x | x << 2) & 0x09249249u;
	return x;
}

// frame / planes of a node along one axis: QuantiseFrame / QuantiseChild of bvh_builder.cpp (same as refit.cu::frameOf / quantise)
__device__ __forceinline__ uint32_t tlasFrame( const float lo, const float hi, float& p )
{
	const double ext = (double)hi - (double)lo;
	int ex = -100;
	p = lo;
	if (ext > 0)
	{
		// frexp( ext / 254.75 ): the exponent field of the double, + 1 for the [0.5, 1) mantissa (no denormals: ext comes from floats)
		ex = ((__double2hiint( ext / 254.75 ) >> 20) & 0x7ff) - 1022;
		if (__hiloint2double( (ex - 1 + 1023) << 20, 0 ) * 254.75 >= ext) ex--;
		if (ex < -100) ex = -100;
		p = (float)((double)lo - __hiloint2double( (ex - 3 + 1023) << 20, 0 ));
	}
	return (uint32_t)(ex + 127);
}
__device__ __forceinline__ void tlasQuantise( const float p, const uint32_t e, const float lo, const float hi, uint32_t& qlo, uint32_t& qhi )
{
	const double inv = __hiloint2double( (127 - (int)e + 1023) << 20, 0 ); // 2^(127 - e)
	const double l = floor( ((double)lo - (double)p) * inv - 1.0 / 64.0 ), h = ceil( ((double)hi - (double)p) * inv + 1.0 / 64.0 );
	qlo = (uint32_t)(l < 0 ? 0 : (l > 255 ? 255 : l)), qhi = (uint32_t)(h < 0 ? 0 : (h > 255 ? 255 : h));
}

struct TlasScratch { float4 *primLo, *primHi, *planLo, *planHi, *wideLo, *wideHi; uint32_t *instOrder; };

__global__ void __launch_bounds__( 1024 ) tlasBuildKernel( const TlasInstanceInput* __restrict__ in, const uint32_t n, const uint32_t P /* power of two >= n */,
	uint4* nodes, const uint32_t nodeBase, const uint32_t maxNodes, float4* __restrict__ instRecs, float4* __restrict__ instShade, uint32_t* info, const TlasScratch sc,
	const float primCost )
{
	extern __shared__ __align__( 16 ) unsigned char smem[];
	// shared-memory carve-up (P elements each unless noted)
	uint64_t* keys = reinterpret_cast<uint64_t*>( smem );
	float4* nodeLo = reinterpret_cast<float4*>( keys + P );
	float4* nodeHi = nodeLo + P;
	uint32_t* sortedPrim = reinterpret_cast<uint32_t*>( nodeHi + P );
	uint32_t *childL = sortedPrim + P, *childR = childL + P, *rangeFirst = childR + P, *rangeCount = rangeFirst + P, *parentInner = rangeCount + P;
	uint32_t *parentLeaf = parentInner + P, *visits = parentLeaf + P, *wideToBinary = visits + P; // wideToBinary: P + 16
	// the collapse plan of the binary nodes (32 bytes each).  In shared memory: the bottom-up pass is a chain as long as the tree is deep, walked by
	// ONE thread at the end, and with the plans in global scratch every link of it paid two L2 round trips and a device-wide fence (two thirds of
	// the kernel's 269 us at 1000 instances, ncu source view of round 2)
	float4* planLo = reinterpret_cast<float4*>( wideToBinary + P + 16 );
	float4* planHi = planLo + P;
	__shared__ int sBounds[6];
	__shared__ uint32_t sCounters[4], sLevel[66], sLevels;
	const uint32_t tid = threadIdx.x, T = blockDim.x;
	// phase clock of thread 0 (SM cycles since the start), kept in info[4..12]: boxes, keys, sort, radix tree, bottom-up, collapse, planes, records
	const long long clock0 = clock64();
	int phase = 4;
	auto stamp = [&]() { if (tid == 0 && phase < 13) info[phase] = (uint32_t)(clock64() - clock0); phase++; };
	if (tid < 6) sBounds[tid] = tid < 3 ? 0x7fffffff : (int)0x80000000;
	if (tid < 4) sCounters[tid] = tid == 0 ? 1u : 0u; // wide node 0 = the root is taken
	__syncthreads();
	// 1. world boxes: all 8 transformed corners of the mesh box (the CPU reference transforms only two, bvh.cpp:579-587: exact for translations,
	//    too small under rotation) plus an ulp-scale pad, since the kernels transform the RAY instead of the box; centroid bounds
	for (uint32_t i = tid; i < n; i += T)
	{
		// (the 176-byte input record as eleven 16-byte loads: three rows of the transform, the two bounds, the flag word -- read field by
		// field, with the records 176 bytes apart, every scalar load of the warp touched its own sector)
		const float4* I4 = reinterpret_cast<const float4*>( in + i );
		const float4 t0 = __ldg( I4 ), t1 = __ldg( I4 + 1 ), t2 = __ldg( I4 + 2 ), blo = __ldg( I4 + 8 ), bhi = __ldg( I4 + 9 );
		const bool empty = __float_as_uint( __ldg( I4 + 10 ).z ) != 0;
		const float c[12] = { t0.x, t0.y, t0.z, t0.w, t1.x, t1.y, t1.z, t1.w, t2.x, t2.y, t2.z, t2.w };
		float lo[3] = { 3.0e38f, 3.0e38f, 3.0e38f }, hi[3] = { -3.0e38f, -3.0e38f, -3.0e38f };
		if (empty) { lo[0] = hi[0] = c[3], lo[1] = hi[1] = c[7], lo[2] = hi[2] = c[11]; }
		else
		{
#pragma unroll
			for (int k = 0; k < 8; k++)
			{
				const float x = (k & 1) ? bhi.x : blo.x, y = (k & 2) ? bhi.y : blo.y, z = (k & 4) ? bhi.z : blo.z;
				const float p[3] = { c[0] * x + c[1] * y + c[2] * z + c[3], c[4] * x + c[5] * y + c[6] * z + c[7], c[8] * x + c[9] * y + c[10] * z + c[11] };
				for (int a = 0; a < 3; a++) lo[a] = fminf( lo[a], p[a] ), hi[a] = fmaxf( hi[a], p[a] );
			}
			for (int a = 0; a < 3; a++)
			{
				const float pad = 4e-7f * fmaxf( fabsf( lo[a] ), fabsf( hi[a] ) ) + 1e-30f;
				lo[a] -= pad, hi[a] += pad;
			}
		}
		sc.primLo[i] = make_float4( lo[0], lo[1], lo[2], 0 ), sc.primHi[i] = make_float4( hi[0], hi[1], hi[2], 0 );
		for (int a = 0; a < 3; a++)
		{
			const int o = orderedInt( 0.5f * (lo[a] + hi[a]) );
			atomicMin( &sBounds[a], o ), atomicMax( &sBounds[3 + a], o );
		}
	}
	__syncthreads();
	stamp();
	// 2. keys: 30-bit Morton code of the centroid, the instance index below it (unique keys: no ties for the radix tree); padding sorts last
	for (uint32_t i = tid; i < P; i += T)
	{
		uint64_t key = ~0ull;
		if (i < n)
		{
			const float4 lo = sc.primLo[i], hi = sc.primHi[i];
			const float c[3] = { 0.5f * (lo.x + hi.x), 0.5f * (lo.y + hi.y), 0.5f * (lo.z + hi.z) };
			uint32_t q[3];
			for (int a = 0; a < 3; a++)
			{
				const float mn = orderedFloat( sBounds[a] ), ext = orderedFloat( sBounds[3 + a] ) - mn;
				const float t = fminf( fmaxf( ext > 0 ? (c[a] - mn) / ext : 0.f, 0.f ), 1.f );
				const uint32_t g = (uint32_t)(t * 1023.0f);
				q[a] = g > 1023u ? 1023u : g;
			}
			key = ((uint64_t)(spread10( q[0] ) << 2 | spread10( q[1] ) << 1 | spread10( q[2] )) << 32) | i;
		}
		keys[i] = key;
	}
	__syncthreads();
	stamp();
	// 3. bitonic sort of the P keys
	for (uint32_t k = 2; k <= P; k <<= 1)
		for (uint32_t j = k >> 1; j > 0; j >>= 1)
		{
			for (uint32_t i = tid; i < P; i += T)
			{
				const uint32_t partner = i ^ j;
				if (partner > i)
				{
					const uint64_t a = keys[i], b = keys[partner];
					if (((i & k) == 0) == (a > b)) keys[i] = b, keys[partner] = a;
				}
			}
			__syncthreads();
		}
	for (uint32_t i = tid; i < n; i += T) sortedPrim[i] = (uint32_t)keys[i], visits[i] = 0;
	__syncthreads();
	stamp();
	if (n == 1)
	{
		// a single instance: one node with one leaf slot
		if (tid == 0)
		{
			sc.instOrder[0] = sortedPrim[0];
			nodes[0] = make_uint4( 0, 0, 0, 0 ), nodes[1] = make_uint4( nodeBase, 0, (1u << 5), 0 ), nodes[2] = nodes[3] = nodes[4] = make_uint4( 0, 0, 0, 0 );
			sCounters[0] = 1, sLevel[0] = 0, sLevel[1] = 1, sLevels = 1;
		}
	}
	else
	{
		// 4. Karras radix tree over the sorted keys (no duplicates)
		for (uint32_t ii = tid; ii + 1 < n; ii += T)
		{
			const int i = (int)ii, nn = (int)n;
			auto delta = [&]( const int a, const int b ) -> int { return (b < 0 || b >= nn) ? -1 : __clzll( (long long)(keys[a] ^ keys[b]) ); };
			const int d = delta( i, i + 1 ) - delta( i, i - 1 ) >= 0 ? 1 : -1;
			const int dmin = delta( i, i - d );
			int lmax = 2;
			while (delta( i, i + lmax * d ) > dmin) lmax *= 2;
			int l = 0;
			for (int t = lmax / 2; t >= 1; t /= 2) if (delta( i, i + (l + t) * d ) > dmin) l += t;
			const int j = i + l * d, dnode = delta( i, j );
			int s = 0;
			for (int div = 2, t = (l + div - 1) / div; ; div *= 2, t = (l + div - 1) / div)
			{
				if (delta( i, i + (s + t) * d ) > dnode) s += t;
				if (t <= 1) break;
			}
			const int gamma = i + s * d + min( d, 0 ), first = min( i, j ), last = max( i, j );
			const uint32_t L = first == gamma ? (LEAF | (uint32_t)gamma) : (uint32_t)gamma;
			const uint32_t R = last == gamma + 1 ? (LEAF | (uint32_t)(gamma + 1)) : (uint32_t)(gamma + 1);
			childL[i] = L, childR[i] = R, rangeFirst[i] = (uint32_t)first, rangeCount[i] = (uint32_t)(last - first + 1);
			if (L & LEAF) parentLeaf[gamma] = ii; else parentInner[gamma] = ii;
			if (R & LEAF) parentLeaf[gamma + 1] = ii; else parentInner[gamma + 1] = ii;
			if (i == 0) parentInner[0] = 0xffffffffu;
		}
		__syncthreads();
		stamp();
		// 5. bottom-up: boxes of the binary nodes and the collapse plan (leaf slots hold ONE instance)
		for (uint32_t i = tid; i < n; i += T)
		{
			uint32_t node = parentLeaf[i];
			while (node != 0xffffffffu)
			{
				__threadfence_block();
				if (atomicAdd( &visits[node], 1u ) == 0) break; // the sibling subtree is not done yet
				__threadfence_block();
				float4 lo[2], hi[2];
				float cl[7], cr[7], cost[7];
				const uint32_t c[2] = { childL[node], childR[node] };
				for (int k = 0; k < 2; k++)
				{
					float* cc = k ? cr : cl;
					if (c[k] & LEAF)
					{
						const uint32_t p = sortedPrim[c[k] & ~LEAF];
						lo[k] = sc.primLo[p], hi[k] = sc.primHi[p];
						const float x = hi[k].x - lo[k].x, y = hi[k].y - lo[k].y, z = hi[k].z - lo[k].z, v = (x * y + y * z + z * x) * primCost;
						for (int q = 0; q < 7; q++) cc[q] = v;
					}
					else
					{
						const volatile float *vl = reinterpret_cast<const volatile float*>( &nodeLo[c[k]] ), *vh = reinterpret_cast<const volatile float*>( &nodeHi[c[k]] );
						lo[k] = make_float4( vl[0], vl[1], vl[2], 0 ), hi[k] = make_float4( vh[0], vh[1], vh[2], 0 );
						const volatile float *pa = reinterpret_cast<const volatile float*>( &planLo[c[k]] ), *pb = reinterpret_cast<const volatile float*>( &planHi[c[k]] );
						const float4 a = make_float4( pa[0], pa[1], pa[2], pa[3] ), b = make_float4( pb[0], pb[1], pb[2], 0 );
						cc[0] = a.x, cc[1] = a.y, cc[2] = a.z, cc[3] = a.w, cc[4] = b.x, cc[5] = b.y, cc[6] = b.z;
					}
				}
				const float4 nl = make_float4( fminf( lo[0].x, lo[1].x ), fminf( lo[0].y, lo[1].y ), fminf( lo[0].z, lo[1].z ), 0 );
				const float4 nh = make_float4( fmaxf( hi[0].x, hi[1].x ), fmaxf( hi[0].y, hi[1].y ), fmaxf( hi[0].z, hi[1].z ), 0 );
				nodeLo[node] = nl, nodeHi[node] = nh;
				const float ex = nh.x - nl.x, ey = nh.y - nl.y, ez = nh.z - nl.z;
				uint32_t bits;
				mergePlan( cl, cr, ex * ey + ey * ez + ez * ex, rangeCount[node], 1, primCost, cost, bits );
				planLo[node] = make_float4( cost[0], cost[1], cost[2], cost[3] );
				planHi[node] = make_float4( cost[4], cost[5], cost[6], __uint_as_float( bits ) );
				node = parentInner[node]; // (the block-level fence at the top of the loop orders the stores before the arrival)
			}
		}
		__syncthreads();
		stamp();
		// 6. collapse to 8-wide nodes, level by level
		CollapseArgs a;
		a.sortedPrim = sortedPrim, a.childL = childL, a.childR = childR, a.rangeCount = rangeCount;
		a.primLo = sc.primLo, a.primHi = sc.primHi, a.nodeLo = nodeLo, a.nodeHi = nodeHi, a.planHi = planHi;
		a.wideToBinary = wideToBinary, a.nodes = nodes, a.primOut = sc.instOrder, a.primOutStride = 1, a.counters = sCounters;
		a.nodeBase = nodeBase, a.triBase = 0, a.maxNodes = maxNodes;
		a.levelTotals = nullptr, a.levelBegin = a.levelChildBase = a.levelPrimBase = 0;
		if (tid == 0) wideToBinary[0] = 0, sLevel[0] = 0, sLevels = 0;
		__syncthreads();
		uint32_t begin = 0, end = 1;
		while (begin < end)
		{
#if B200_USE_GROUP_COLLAPSE
			for (uint32_t w = begin + tid / 8; w < end; w += T / 8) collapseNodeGroup( a, w, tid & 7u, 0xffu << ((tid & 31u) - (tid & 7u)) );
#else
			for (uint32_t w = begin + tid; w < end; w += T) collapseNode( a, w );
#endif
			__syncthreads();
			if (tid == 0) sLevel[++sLevels] = end;
			begin = end, end = min( sCounters[0], maxNodes );
			__syncthreads();
			if (sLevels >= 64) break;
		}
	}
	__syncthreads();
	if (n == 1) phase = 9; else stamp();
	// 7. boxes, frames and quantised planes, deepest level first (a leaf slot's box is its instance's world box).  Eight lanes per node,
	//    lane s owns child slot s (as refit.cu::refitSlot does): the node box is an 8-lane shuffle reduction, every lane quantises its own child
	const uint32_t nodeCount = sCounters[0] < maxNodes ? sCounters[0] : maxNodes, levels = sLevels;
	for (int level = (int)levels - 1; level >= 0; level--)
	{
		const uint32_t begin = sLevel[level], slots = (sLevel[level + 1] - begin) * 8;
		for (uint32_t i = tid; i < ((slots + T - 1) / T) * T; i += T) // (whole warps iterate: the shuffles below need every lane of a group's warp)
		{
			const bool valid = i < slots;
			const uint32_t w = begin + (valid ? i >> 3 : 0), slot = i & 7;
			uint4* np = nodes + (size_t)w * 5;
			const uint4 n0 = np[0], n1 = np[1];
			const uint32_t imask = n0.w >> 24;
			const uint32_t meta = ((slot < 4 ? n1.z : n1.w) >> (8 * (slot & 3))) & 0xffu;
			float clo[3] = { 3.0e38f, 3.0e38f, 3.0e38f }, chi[3] = { -3.0e38f, -3.0e38f, -3.0e38f };
			if (valid && meta)
			{
				float4 l, h;
				if (imask & (1u << slot))
				{
					const uint32_t child = n1.x + __popc( imask & ((1u << slot) - 1u) ) - nodeBase;
					l = __ldcg( sc.wideLo + child ), h = __ldcg( sc.wideHi + child );
				}
				else
				{
					const uint32_t inst = __ldcg( sc.instOrder + n1.y + (meta & 31u) );
					l = sc.primLo[inst], h = sc.primHi[inst];
				}
				clo[0] = l.x, clo[1] = l.y, clo[2] = l.z, chi[0] = h.x, chi[1] = h.y, chi[2] = h.z;
			}
			uint32_t e[3], ql[3] = { 0, 0, 0 }, qh[3] = { 0, 0, 0 };
			float fp[3], nlo[3], nhi[3];
			for (int c = 0; c < 3; c++)
			{
				float lo = clo[c], hi = chi[c];
				for (int d = 1; d < 8; d <<= 1) lo = fminf( lo, __shfl_xor_sync( 0xffffffffu, lo, d ) ), hi = fmaxf( hi, __shfl_xor_sync( 0xffffffffu, hi, d ) );
				nlo[c] = lo, nhi[c] = hi;
				e[c] = tlasFrame( lo, hi, fp[c] );
				if (valid && meta) tlasQuantise( fp[c], e[c], clo[c], chi[c], ql[c], qh[c] );
			}
			if (!valid) continue;
			uint8_t* q = reinterpret_cast<uint8_t*>( np + 2 ); // qlox[8] qloy[8] qloz[8] qhix[8] qhiy[8] qhiz[8]
			q[slot] = (uint8_t)ql[0], q[8 + slot] = (uint8_t)ql[1], q[16 + slot] = (uint8_t)ql[2];
			q[24 + slot] = (uint8_t)qh[0], q[32 + slot] = (uint8_t)qh[1], q[40 + slot] = (uint8_t)qh[2];
			if (slot == 0)
			{
				np[0] = make_uint4( __float_as_uint( fp[0] ), __float_as_uint( fp[1] ), __float_as_uint( fp[2] ), e[0] | (e[1] << 8) | (e[2] << 16) | (imask << 24) );
				__stcg( sc.wideLo + w, make_float4( nlo[0], nlo[1], nlo[2], 0 ) ), __stcg( sc.wideHi + w, make_float4( nhi[0], nhi[1], nhi[2], 0 ) );
			}
		}
		__threadfence();
		__syncthreads();
	}
	stamp();
	// 8. instance records in leaf order (wide_bvh.h InstanceRec) and the shading-time table by instance id (rendercore.cpp)
	for (uint32_t j = tid; j < n; j += T)
	{
		const uint32_t id = __ldcg( sc.instOrder + j );
		const float4* I4 = reinterpret_cast<const float4*>( in + id ); // 16-byte loads and stores throughout (see step 1)
		float4* r = instRecs + (size_t)j * 6;
#pragma unroll
		for (int k = 0; k < 4; k++) r[k] = __ldg( I4 + 4 + k ); // inv
		const float4 word = __ldg( I4 + 10 );                    // meshId, flags, empty, -
		r[4] = make_float4( __ldg( I4 + 8 ).w /* blasRoot */, __uint_as_float( id ), word.x, word.y );
		r[5] = make_float4( 0, 0, 0, 0 );
		const float4* S4 = reinterpret_cast<const float4*>( in + j );
		float4* sh = instShade + (size_t)j * 4;
#pragma unroll
		for (int k = 0; k < 3; k++) sh[k] = __ldg( S4 + k );   // transform rows
		sh[3] = make_float4( __ldg( S4 + 9 ).w /* triBase */, __ldg( S4 + 10 ).x /* meshId */, 0, 0 );
	}
	__syncthreads();
	stamp();
	if (tid == 0) info[0] = nodeCount, info[1] = levels, info[2] = (sCounters[2] || sCounters[0] > maxNodes || levels >= 64) ? 1u : 0u, info[3] = 0;
}

} // namespace

// Builds the wide-BVH topology of `T` triangles (dVerts = float4[3T], device) into nodes[0 .. nodeCount) / tris[0 .. T) (pool pointers
// already offset to this mesh's base; indices stored in the nodes are absolute: + nodeBase / + triBase).  levelStart (relative,
// size levels + 1) and nodeCount are returned for the refit, which the caller runs next.  T >= 4.
#ifndef B200_PLOC_RADIUS
#define B200_PLOC_RADIUS 16
#endif
cudaError_t GpuBuildTopology( const float4* dVerts, uint32_t T, uint4* nodes, float4* tris, uint32_t nodeBase, uint32_t triBase, uint32_t maxNodes,
	cudaStream_t s, std::vector<uint32_t>& levelStart, uint32_t& nodeCount, int builder, void** keptScratch, size_t* keptBytes )
{
	cudaError_t e;
	char* scratch = nullptr;
	auto align = []( size_t x ) { return (x + 255) & ~(size_t)255; };
	size_t sortBytes = 0;
	cub::DeviceRadixSort::SortPairs( nullptr, sortBytes, (const uint64_t*)nullptr, (uint64_t*)nullptr, (const uint32_t*)nullptr, (uint32_t*)nullptr, (int)T, 0, 63, s );
	// scratch layout
	size_t off = 0;
	auto take = [&]( size_t bytes ) { const size_t o = off; off += align( bytes ); return o; };
	const size_t oPrimLo = take( (size_t)T * 16 ), oPrimHi = take( (size_t)T * 16 ), oKeys = take( (size_t)T * 8 ), oKeys2 = take( (size_t)T * 8 );
	const size_t oVals = take( (size_t)T * 4 ), oVals2 = take( (size_t)T * 4 ), oSort = take( sortBytes );
	const size_t oChildL = take( (size_t)T * 4 ), oChildR = take( (size_t)T * 4 ), oFirst = take( (size_t)T * 4 ), oCount = take( (size_t)T * 4 );
	const size_t oParI = take( (size_t)T * 4 ), oParL = take( (size_t)T * 4 ), oVisits = take( (size_t)T * 4 );
	const size_t oNodeLo = take( (size_t)T * 16 ), oNodeHi = take( (size_t)T * 16 ), oPlanLo = take( (size_t)T * 16 ), oPlanHi = take( (size_t)T * 16 );
	const size_t oW2B = take( (size_t)maxNodes * 4 + 64 ), oSmall = take( 256 );
	// per-level claims of the collapse and their running sum (children << 32 | primitives per wide node)
	size_t totalBytes = 0;
	cub::DeviceScan::ExclusiveSum( nullptr, totalBytes, (const unsigned long long*)nullptr, (unsigned long long*)nullptr, (int)maxNodes, s );
	const size_t oClaims = take( (size_t)maxNodes * 8 ), oTotals = take( (size_t)maxNodes * 8 ), oTotalScan = take( totalBytes );
	// the agglomerative builder: two sets of clusters (reference + box), nearest neighbours, survivor flags and their running sum
	const bool ploc = builder == 2;
	size_t scanBytes = 0;
	if (ploc) cub::DeviceScan::ExclusiveSum( nullptr, scanBytes, (const uint32_t*)nullptr, (uint32_t*)nullptr, (int)T, s );
	size_t oCRef[2] = { 0, 0 }, oCLo[2] = { 0, 0 }, oCHi[2] = { 0, 0 }, oNearest = 0, oFlag = 0, oOffset = 0, oScan = 0;
	if (ploc)
	{
		for (int k = 0; k < 2; k++) oCRef[k] = take( (size_t)T * 4 ), oCLo[k] = take( (size_t)T * 16 ), oCHi[k] = take( (size_t)T * 16 );
		oNearest = take( (size_t)T * 4 ), oFlag = take( (size_t)T * 4 ), oOffset = take( (size_t)T * 4 ), oScan = take( scanBytes );
	}
	size_t oSpFlag = 0, oSpRef = 0, oSpLast = 0, oItRef = 0, oItFirst = 0, oItLast = 0, oItLo = 0, oItHi = 0, oItBefore = 0;
	if (builder == 3)
	{
		cub::DeviceScan::ExclusiveSum( nullptr, scanBytes, (const uint32_t*)nullptr, (uint32_t*)nullptr, (int)T, s );
		oSpFlag = take( (size_t)T * 4 ), oSpRef = take( (size_t)T * 4 ), oSpLast = take( (size_t)T * 4 ), oItRef = take( (size_t)T * 4 ), oItFirst = take( (size_t)T * 4 );
		oItLast = take( (size_t)T * 4 ), oItLo = take( (size_t)T * 16 ), oItHi = take( (size_t)T * 16 ), oItBefore = take( (size_t)T * 4 ), oScan = take( scanBytes );
	}
	// builder 4: the large nodes of two levels (at most T / SAHB_CLUSTER each), their bins and splits, the list of small subtrees
	const uint32_t topCap = T / SAHB_CLUSTER + 2;
	size_t oTopNodes[2] = { 0, 0 }, oTopBins = 0, oTopSplit = 0, oTopSmall = 0;
	if (builder == 4)
	{
		cub::DeviceScan::ExclusiveSum( nullptr, scanBytes, (const uint32_t*)nullptr, (uint32_t*)nullptr, (int)T, s );
		for (int k = 0; k < 2; k++) oTopNodes[k] = take( (size_t)topCap * 36 );
		oTopBins = take( (size_t)topCap * 3 * TOP_BINS * 28 ), oTopSplit = take( (size_t)topCap * 16 ), oTopSmall = take( ((size_t)T / 2 + 2) * 16 ), oScan = take( scanBytes );
	}
	static const bool timing = getenv( "B200_BUILD_TIMING" ) != nullptr;
	auto clockNow = [] { return std::chrono::duration<double, std::milli>( std::chrono::steady_clock::now().time_since_epoch() ).count(); };
	double tMark = clockNow();
	auto mark = [&]( const char* what ) { if (!timing) return; cudaStreamSynchronize( s ); const double t = clockNow(); fprintf( stderr, "  topology: %s %.2f ms\n", what, t - tMark ); tMark = t; };
	// the caller may own the scratch (a core that rebuilds its meshes keeps it between builds: allocating and freeing 190 MB per million
	// triangles was 10 of the 16 ms of a build); otherwise it lives for this call
	const bool kept = keptScratch && keptBytes;
	if (kept)
	{
		if (*keptBytes < off)
		{
			if (*keptScratch) cudaStreamSynchronize( s ), cudaFree( *keptScratch );
			*keptScratch = nullptr, *keptBytes = 0;
			if ((e = cudaMalloc( keptScratch, off )) != cudaSuccess) return e;
			*keptBytes = off;
		}
		scratch = (char*)*keptScratch;
	}
	else if ((e = cudaMallocAsync( (void**)&scratch, off, s )) != cudaSuccess) return e;
	if (timing) fprintf( stderr, "  topology: %.1f MB of scratch\n", off / 1048576.0 );
	mark( "scratch allocation" );
	float4 *primLo = (float4*)(scratch + oPrimLo), *primHi = (float4*)(scratch + oPrimHi), *nodeLo = (float4*)(scratch + oNodeLo), *nodeHi = (float4*)(scratch + oNodeHi);
	float4 *planLo = (float4*)(scratch + oPlanLo), *planHi = (float4*)(scratch + oPlanHi);
	uint64_t *keys = (uint64_t*)(scratch + oKeys), *keys2 = (uint64_t*)(scratch + oKeys2);
	uint32_t *vals = (uint32_t*)(scratch + oVals), *vals2 = (uint32_t*)(scratch + oVals2);
	uint32_t *childL = (uint32_t*)(scratch + oChildL), *childR = (uint32_t*)(scratch + oChildR), *first = (uint32_t*)(scratch + oFirst), *count = (uint32_t*)(scratch + oCount);
	uint32_t *parI = (uint32_t*)(scratch + oParI), *parL = (uint32_t*)(scratch + oParL), *visits = (uint32_t*)(scratch + oVisits), *w2b = (uint32_t*)(scratch + oW2B);
	int* cbounds = (int*)(scratch + oSmall);
	uint32_t* counters = (uint32_t*)(scratch + oSmall + 64);
	auto done = [&]( cudaError_t err ) { if (!kept) cudaFreeAsync( scratch, s ); return err; };

	const int initBounds[6] = { 0x7fffffff, 0x7fffffff, 0x7fffffff, (int)0x80000000, (int)0x80000000, (int)0x80000000 };
	const uint32_t initCounters[4] = { 1, 0, 0, 0 }; // wide node 0 = the root is taken
	const uint32_t rootBinary = 0;
	if ((e = cudaMemcpyAsync( cbounds, initBounds, sizeof( initBounds ), cudaMemcpyHostToDevice, s )) != cudaSuccess) return done( e );
	if ((e = cudaMemcpyAsync( counters, initCounters, sizeof( initCounters ), cudaMemcpyHostToDevice, s )) != cudaSuccess) return done( e );
	if ((e = cudaMemcpyAsync( w2b, &rootBinary, 4, cudaMemcpyHostToDevice, s )) != cudaSuccess) return done( e );
	if ((e = cudaMemsetAsync( visits, 0, (size_t)T * 4, s )) != cudaSuccess) return done( e );
	const uint32_t blocks = (T + 255) / 256;
	primBoundsKernel<<<blocks, 256, 0, s>>>( dVerts, T, primLo, primHi, cbounds );
	if (builder == 4)
	{
		// binned SAH from the top (see 3d): no sort, no radix tree.  order / nodeOf: two copies each (the key arrays serve as scratch)
		uint32_t *order[2] = { vals, vals2 }, *nodeOf[2] = { (uint32_t*)keys, (uint32_t*)keys + T }, *flag = (uint32_t*)keys2, *leftBefore = (uint32_t*)keys2 + T;
		TopNodes act[2];
		for (int k = 0; k < 2; k++)
		{
			char* b = scratch + oTopNodes[k];
			act[k].id = (uint32_t*)b, act[k].begin = (uint32_t*)b + topCap, act[k].count = (uint32_t*)b + 2 * topCap, act[k].cb = (int*)b + 3 * topCap;
		}
		TopBins bins;
		bins.count = (int*)(scratch + oTopBins), bins.lo = bins.count + (size_t)topCap * 3 * TOP_BINS, bins.hi = bins.lo + (size_t)topCap * 3 * TOP_BINS * 3;
		uint4 *split = (uint4*)(scratch + oTopSplit), *smallRoots = (uint4*)(scratch + oTopSmall);
		uint32_t *nextCount = counters + 8, *smallCount = counters + 9, *nodeCounter = counters + 10;
		const uint32_t init[3] = { 0, 0, 1 }; // (node 0 is the root)
		if ((e = cudaMemcpyAsync( nextCount, init, 12, cudaMemcpyHostToDevice, s )) != cudaSuccess) return done( e );
		topInitKernel<<<blocks, 256, 0, s>>>( T, order[0], nodeOf[0], act[0], cbounds, parI );
		int cur = 0;
		uint32_t nAct = 1, nSmall = 0;
		if (T <= SAHB_CLUSTER)
		{
			const uint4 whole = make_uint4( 0, 0, T, 0 );
			if ((e = cudaMemcpyAsync( smallRoots, &whole, 16, cudaMemcpyHostToDevice, s )) != cudaSuccess) return done( e );
			nAct = 0, nSmall = 1;
			if ((e = cudaMemcpyAsync( smallCount, &nSmall, 4, cudaMemcpyHostToDevice, s )) != cudaSuccess) return done( e );
		}
		for (int level = 0; nAct; level++)
		{
			if (level > 4096) return done( cudaErrorUnknown ); // (every level splits every large node into two non-empty parts)
			const uint32_t clearThreads = max( nAct * 3 * TOP_BINS, min( 2 * nAct, topCap ) * 6 );
			topClearBinsKernel<<<(clearThreads + 255) / 256, 256, 0, s>>>( nAct, bins, act[cur ^ 1].cb, min( 2 * nAct, topCap ) );
			topBinKernel<<<blocks, 256, 0, s>>>( T, order[cur], nodeOf[cur], act[cur], primLo, primHi, bins );
			topSplitKernel<<<(nAct * 32 + 127) / 128, 128, 0, s>>>( nAct, act[cur], bins, split, act[cur ^ 1], nextCount, smallRoots, smallCount, nodeCounter, childL, childR, count, parI, parL );
			topFlagKernel<<<blocks, 256, 0, s>>>( T, order[cur], nodeOf[cur], act[cur], split, primLo, primHi, flag );
			if ((e = cub::DeviceScan::ExclusiveSum( scratch + oScan, scanBytes, flag, leftBefore, (int)T, s )) != cudaSuccess) return done( e );
			topScatterKernel<<<blocks, 256, 0, s>>>( T, order[cur], nodeOf[cur], act[cur], split, primLo, primHi, leftBefore, order[cur ^ 1], nodeOf[cur ^ 1], act[cur ^ 1].cb );
			uint32_t h[2];
			if ((e = cudaMemcpyAsync( h, nextCount, 8, cudaMemcpyDeviceToHost, s )) != cudaSuccess) return done( e );
			if ((e = cudaMemsetAsync( nextCount, 0, 4, s )) != cudaSuccess) return done( e );
			if ((e = cudaStreamSynchronize( s )) != cudaSuccess) return done( e );
			if (getenv( "B200_BUILD_TIMING" )) fprintf( stderr, "top-down level %d: %u large nodes -> %u, small subtrees so far %u\n", level, nAct, h[0], h[1] );
			nAct = h[0], nSmall = h[1], cur ^= 1;
			if (nAct > topCap) return done( cudaErrorUnknown );
		}
		mark( "large nodes, level by level" );
		if (nSmall)
		{
			static bool allowed = false;
			const size_t smem = sizeof( SahbWarp ) * SAHB_WARPS;
			if (!allowed)
			{
				if ((e = cudaFuncSetAttribute( sahbBuildKernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem )) != cudaSuccess) return done( e );
				allowed = true;
			}
			sahbBuildKernel<false><<<(nSmall + SAHB_WARPS - 1) / SAHB_WARPS, SAHB_WARPS * 32, smem, s>>>( smallRoots, smallCount, SahbItems(), SahbSparse(), childL, childR, count, parI, parL,
				nodeLo, nodeHi, order[cur], primLo, primHi, nodeCounter );
		}
		if (order[cur] != vals2) { uint32_t* t = vals; vals = vals2, vals2 = t; } // (what follows calls the final order vals2)
		mark( "small subtrees" );
		binaryBoxesKernel<<<blocks, 256, 0, s>>>( vals2, primLo, primHi, (int)T, childL, childR, count, parI, parL, nodeLo, nodeHi, planLo, planHi, g_sahPrimCost, visits );
	}
	else
	{
	mortonKernel<<<blocks, 256, 0, s>>>( primLo, primHi, T, cbounds, keys, vals );
	if ((e = cub::DeviceRadixSort::SortPairs( scratch + oSort, sortBytes, keys, keys2, vals, vals2, (int)T, 0, 63, s )) != cudaSuccess) return done( e );
	}
	if (builder == 4) {}
	else if (!ploc)
	{
		radixTreeKernel<<<blocks, 256, 0, s>>>( keys2, (int)T, childL, childR, first, count, parI, parL );
		if (builder == 3)
		{
			// SAH inside the radix tree's cells, level by level (see 3c)
			SahbSparse sparse = { (uint32_t*)(scratch + oSpFlag), (uint32_t*)(scratch + oSpRef), (uint32_t*)(scratch + oSpLast) };
			SahbItems items = { (uint32_t*)(scratch + oItRef), (uint32_t*)(scratch + oItFirst), (uint32_t*)(scratch + oItLast), (float4*)(scratch + oItLo), (float4*)(scratch + oItHi) };
			uint32_t *before = (uint32_t*)(scratch + oItBefore), *rootCount = counters + 8;
			uint4* roots = (uint4*)keys; // (the first key array is free after the sort: T / 2 roots of 16 bytes)
			static bool allowed = false;
			const size_t smem = sizeof( SahbWarp ) * SAHB_WARPS;
			if (!allowed)
			{
				if ((e = cudaFuncSetAttribute( sahbBuildKernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem )) != cudaSuccess) return done( e );
				allowed = true;
			}
			sahbInitKernel<<<blocks, 256, 0, s>>>( T, sparse );
			// (levels above the first: tools/ab_sahb.sh -- on the sphere field they trade primary rays for shadow rays, -3.5 % / +2.8 %; the default is 1)
			for (int level = 0; level < g_sahLevels; level++)
			{
				if (level > 32) return done( cudaErrorUnknown );
				if ((e = cub::DeviceScan::ExclusiveSum( scratch + oScan, scanBytes, sparse.flag, before, (int)T, s )) != cudaSuccess) return done( e );
				if ((e = cudaMemsetAsync( rootCount, 0, 4, s )) != cudaSuccess) return done( e );
				sahbCompactKernel<<<blocks, 256, 0, s>>>( T, sparse, before, vals2, primLo, primHi, nodeLo, nodeHi, items );
				sahbRootsKernel<<<blocks, 256, 0, s>>>( T, sparse, before, items, first, count, parI, roots, rootCount, (uint32_t)SAHB_CLUSTER );
				uint32_t h[2]; // items of this level (the last position's running sum + its flag: position T - 1 always belongs to the last item, whose start is counted before it unless it is a single), roots
				uint32_t lastBefore = 0, lastFlag = 0, nRoots = 0;
				if ((e = cudaMemcpyAsync( &lastBefore, before + (T - 1), 4, cudaMemcpyDeviceToHost, s )) != cudaSuccess) return done( e );
				if ((e = cudaMemcpyAsync( &lastFlag, sparse.flag + (T - 1), 4, cudaMemcpyDeviceToHost, s )) != cudaSuccess) return done( e );
				if ((e = cudaMemcpyAsync( &nRoots, rootCount, 4, cudaMemcpyDeviceToHost, s )) != cudaSuccess) return done( e );
				if ((e = cudaStreamSynchronize( s )) != cudaSuccess) return done( e );
				h[0] = lastBefore + lastFlag, h[1] = nRoots;
				if (h[0] <= 1 || !h[1]) break; // everything is inside the root's subtree
				sahbBuildKernel<true><<<(h[1] + SAHB_WARPS - 1) / SAHB_WARPS, SAHB_WARPS * 32, smem, s>>>( roots, rootCount, items, sparse, childL, childR, count, parI, parL, nodeLo, nodeHi );
			}
		}
		binaryBoxesKernel<<<blocks, 256, 0, s>>>( vals2, primLo, primHi, (int)T, childL, childR, count, parI, parL, nodeLo, nodeHi, planLo, planHi, g_sahPrimCost, visits );
	}
	else
	{
		PlocClusters cl[2];
		for (int k = 0; k < 2; k++) cl[k].ref = (uint32_t*)(scratch + oCRef[k]), cl[k].lo = (float4*)(scratch + oCLo[k]), cl[k].hi = (float4*)(scratch + oCHi[k]);
		uint32_t *nearest = (uint32_t*)(scratch + oNearest), *flag = (uint32_t*)(scratch + oFlag), *offset = (uint32_t*)(scratch + oOffset);
		uint32_t *nPtr = counters + 8, *mergeCounter = counters + 9; // (the 256-byte block of small values)
		const uint32_t init[2] = { T, 0 };
		if ((e = cudaMemcpyAsync( nPtr, init, 8, cudaMemcpyHostToDevice, s )) != cudaSuccess) return done( e );
		plocInitKernel<<<blocks, 256, 0, s>>>( vals2, primLo, primHi, T, cl[0] );
		uint32_t n = T;
		int cur = 0;
		for (int round = 0; n > 1; round++)
		{
			if (round > 4096) return done( cudaErrorUnknown ); // (every round merges at least its smallest pair)
			const uint32_t b = (n + 255) / 256;
			plocNearestKernel<B200_PLOC_RADIUS><<<b, 256, 0, s>>>( cl[cur], nPtr, nearest );
			plocMergeKernel<<<b, 256, 0, s>>>( cl[cur], nPtr, nearest, flag, mergeCounter, T, vals2, primLo, primHi, childL, childR, count, nodeLo, nodeHi, planLo, planHi, g_sahPrimCost );
			if ((e = cub::DeviceScan::ExclusiveSum( scratch + oScan, scanBytes, flag, offset, (int)n, s )) != cudaSuccess) return done( e );
			plocCompactKernel<<<b, 256, 0, s>>>( cl[cur], cl[cur ^ 1], nPtr, n, flag, offset );
			if ((e = cudaMemcpyAsync( &n, nPtr, 4, cudaMemcpyDeviceToHost, s )) != cudaSuccess) return done( e );
			if ((e = cudaStreamSynchronize( s )) != cudaSuccess) return done( e );
			cur ^= 1;
		}
	}
	mark( "binary tree, boxes and plans" );
	if ((e = cudaGetLastError()) != cudaSuccess) return done( e );
	CollapseArgs a;
	a.sortedPrim = vals2, a.childL = childL, a.childR = childR, a.rangeCount = count;
	a.primLo = primLo, a.primHi = primHi, a.nodeLo = nodeLo, a.nodeHi = nodeHi, a.planHi = planHi;
	a.wideToBinary = w2b, a.nodes = nodes, a.primOut = reinterpret_cast<uint32_t*>( tris ) + 3, a.primOutStride = 12, a.counters = counters;
	a.nodeBase = nodeBase, a.triBase = triBase, a.maxNodes = maxNodes;
	unsigned long long *claims = (unsigned long long*)(scratch + oClaims), *totals = (unsigned long long*)(scratch + oTotals);
	a.levelTotals = B200_USE_GROUP_COLLAPSE ? totals : nullptr, a.levelBegin = 0, a.levelChildBase = 1, a.levelPrimBase = 0;
	levelStart.clear();
	levelStart.push_back( 0 );
	uint32_t begin = 0, end = 1;
	while (begin < end)
	{
		const uint32_t levelBlocks = ((end - begin) * (B200_USE_GROUP_COLLAPSE ? 8 : 1) + 127) / 128;
		if (a.levelTotals)
		{
			// the same bytes every time: each node's place follows from the claims of the nodes before it in the level, not from the order
			// in which the groups reach an atomicAdd.  Hits do not depend on the placement (traverse.cu's tie rule), but a rebuilt mesh and
			// the copies of a mesh on several ranks then share one layout (same cache behaviour, comparable dumps), and siblings of
			// neighbouring parents sit next to each other
			a.levelBegin = begin;
			collapseCountKernel<<<levelBlocks, 128, 0, s>>>( a, begin, end, claims );
			if ((e = cub::DeviceScan::ExclusiveSum( scratch + oTotalScan, totalBytes, claims, totals, (int)(end - begin), s )) != cudaSuccess) return done( e );
		}
		collapseKernel<<<levelBlocks, 128, 0, s>>>( a, begin, end );
		uint32_t h[3];
		if ((e = cudaMemcpyAsync( h, counters, 12, cudaMemcpyDeviceToHost, s )) != cudaSuccess) return done( e );
		if ((e = cudaStreamSynchronize( s )) != cudaSuccess) return done( e );
		if (h[2] || h[0] > maxNodes) return done( cudaErrorMemoryAllocation ); // more wide nodes than reserved
		levelStart.push_back( end );
		begin = end, end = h[0];
		a.levelChildBase = h[0], a.levelPrimBase = h[1];
	}
	nodeCount = end;
	mark( "collapse" );
	return done( cudaSuccess );
}

size_t TlasScratchBytes( uint32_t n )
{
	return (size_t)(n + 16) * (6 * 16 + 4) + 256;
}

cudaError_t LaunchTlasBuild( const TlasInstanceInput* dInstances, uint32_t n, uint4* nodes, uint32_t nodeBase, uint32_t maxNodes, float4* instRecs, float4* instShade,
	uint32_t* info, void* scratch, float primCost, cudaStream_t s )
{
	if (n == 0 || n > MAX_DEVICE_TLAS_INSTANCES) return cudaErrorInvalidValue;
	uint32_t P = 4; // (at least 4: the float4 arrays behind the uint32 ones stay 16-byte aligned)
	while (P < n) P <<= 1;
	const size_t smem = (size_t)P * (8 + 32 + 4 * 9 + 32) + 64; // keys, binary node boxes, nine uint32 arrays (wideToBinary with 16 spare entries), collapse plans
	static size_t allowed = 0;
	cudaError_t e;
	if (smem > allowed)
	{
		if ((e = cudaFuncSetAttribute( tlasBuildKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem )) != cudaSuccess) return e;
		allowed = smem;
	}
	TlasScratch sc;
	char* p = (char*)scratch;
	const size_t stride = (size_t)(n + 16) * 16;
	sc.primLo = (float4*)p, sc.primHi = (float4*)(p + stride), sc.planLo = (float4*)(p + 2 * stride), sc.planHi = (float4*)(p + 3 * stride);
	sc.wideLo = (float4*)(p + 4 * stride), sc.wideHi = (float4*)(p + 5 * stride), sc.instOrder = (uint32_t*)(p + 6 * stride);
	const int threads = P < 64 ? 64 : (P > 1024 ? 1024 : (int)P);
	tlasBuildKernel<<<1, threads, smem, s>>>( dInstances, n, P, nodes, nodeBase, maxNodes, instRecs, instShade, info, sc, primCost );
	return cudaGetLastError();
}

} // namespace b200
