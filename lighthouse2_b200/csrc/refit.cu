// refit.cu -- GPU refit of a mesh's 8-wide BVH after its vertices moved (animated meshes, config C4).
//
// The reference has no refit: RenderCore::SetGeometry does a full SAH rebuild on every update
// (lib/RenderCore_Bart/rendercore.cpp:78 -> bvh.cpp:57); the OptiX cores call optixAccelBuild with
// OPTIX_BUILD_OPERATION_UPDATE (lib/rendercore_optix7/core_mesh.cpp:71-103).  Here the topology built by
// bvh_builder.cpp is kept and only boxes are recomputed:
//   refitLevel, deepest level first (nodes are stored breadth-first, so a level is an index range): eight lanes per node, lane s
//   owns child slot s.  A leaf slot gathers the new vertices of its <= 3 triangles once (original primitive order, float4[3T]),
//   rewrites their leaf-order records {v0|prim, e1, e2} and shading normals, and bounds them; an inner slot reads the child's
//   float box from scratch; an 8-lane shuffle reduction gives the node box; each lane re-quantises its own child against the
//   new frame (double arithmetic, same formulas as QuantiseFrame/QuantiseChild in bvh_builder.cpp).
//   Bytes: per triangle 16 (record .w) + 48 (vertices) read, 48 + 16 written; per node 80 read + 80 written + 32 scratch.
//   refitWalkKernel does the whole tree in ONE launch: an 8-lane group starts at every node without inner children and climbs -- the
//   group that delivers the last child box of a node (one atomic counter per node) goes on with that node.  No launch per level (a
//   1 M-triangle tree has 16 levels of ~6 us launch latency each; C4's 100 K-triangle mesh was nothing but launch latency) and no
//   grid-wide barrier: the critical path is the depth of the tree times one node step.
#include "kernels.h"

namespace b200 {

namespace {

// 2^k as a double, k within the normal range (what ldexp( 1.0, k ) returns, without the library call)
__device__ __forceinline__ double pow2( const int k ) { return __hiloint2double( (k + 1023) << 20, 0 ); }

// frame of a node along one axis: exponent byte and origin (restates QuantiseFrame of bvh_builder.cpp: s/8 of headroom on both sides, so
// that every child plane keeps its guard band)
__device__ __forceinline__ uint32_t frameOf( float lo, float hi, float& p )
{
	const double ext = (double)hi - (double)lo;
	int ex = -100;
	p = lo;
	if (ext > 0)
	{
		// QuantiseFrame's frexp( ext / 254.75 ) and its m == 0.5 correction amount to: the smallest ex with 254.75 * 2^ex >= ext.  With
		// ext = m * 2^E, m in [1, 2) (no denormals here: ext comes from floats) and 254.75 = 1.990234375 * 2^7 that is E - 7 or E - 6,
		// decided by one exact product: no division
		ex = ((__double2hiint( ext ) >> 20) & 0x7ff) - 1023 - 7;
		if (pow2( ex ) * 254.75 < ext) ex++;
		if (ex < -100) ex = -100;
		p = (float)((double)lo - pow2( ex - 3 ));
	}
	return (uint32_t)(ex + 127);
}

__device__ __forceinline__ void quantise( float p, uint32_t e, float lo, float hi, uint32_t& qlo, uint32_t& qhi )
{
	const double guard = 1.0 / 64.0;
	const double inv = pow2( 127 - (int)e );
	double l = floor( ((double)lo - (double)p) * inv - guard ), h = ceil( ((double)hi - (double)p) * inv + guard );
	qlo = (uint32_t)(l < 0 ? 0 : (l > 255 ? 255 : l));
	qhi = (uint32_t)(h < 0 ? 0 : (h > 255 ? 255 : h));
}

// a leaf slot: the lane owns the slot's <= 3 triangles.  Gathers their new vertices ONCE (original primitive order), rewrites the
// leaf-order records {v0|prim, e1, e2} and the shading normal, and bounds them.  (All the primitive indices first, then all the vertices:
// one round trip to memory each for the slot, not one per triangle.)
__device__ __forceinline__ void leafSlotBox( const RefitArgs& a, const uint4 n1, const uint32_t meta, float lo[3], float hi[3] )
{
	const uint32_t cnt = __popc( meta >> 5 ), first = n1.y + (meta & 31u);
	float4* const tp0 = a.tris + (size_t)first * 3;
	uint32_t prims[3], words[3] = { 0, 0, 0 };
	float4 v[3][3];
	// which triangle sits in a record and what its shading record carries beside the normal come from a leaf-order copy made when the
	// tree was placed (8 bytes per triangle, neighbouring lanes read neighbouring entries) -- not from the 48-byte records and the
	// shading records themselves, which would cost a 32-byte sector each for 4 bytes
	const uint2* li = a.leafInfo ? a.leafInfo + (first - a.triBase) : nullptr;
#pragma unroll
	for (uint32_t k = 0; k < 3; k++)
	{
		prims[k] = 0;
		if (k < cnt) { if (li) { const uint2 e = __ldg( li + k ); prims[k] = e.x, words[k] = e.y; } else prims[k] = __float_as_uint( tp0[k * 3].w ); }
	}
#pragma unroll
	for (uint32_t k = 0; k < 3; k++) if (k < cnt)
	{
		const float4* vp = a.verts + (size_t)prims[k] * 3;
		v[k][0] = __ldg( vp ), v[k][1] = __ldg( vp + 1 ), v[k][2] = __ldg( vp + 2 );
	}
#pragma unroll
	for (uint32_t k = 0; k < 3; k++) if (k < cnt)
	{
		float4* tp = tp0 + k * 3;
		const uint32_t prim = prims[k];
		const float4 p0 = v[k][0], p1 = v[k][1], p2 = v[k][2];
		const float e1x = p1.x - p0.x, e1y = p1.y - p0.y, e1z = p1.z - p0.z, e2x = p2.x - p0.x, e2y = p2.y - p0.y, e2z = p2.z - p0.z;
		tp[0] = make_float4( p0.x, p0.y, p0.z, __uint_as_float( prim ) );
		tp[1] = make_float4( e1x, e1y, e1z, 0.f );
		tp[2] = make_float4( e2x, e2y, e2z, 0.f );
		if (a.shade)
		{
			float nx = e1y * e2z - e1z * e2y, ny = e1z * e2x - e1x * e2z, nz = e1x * e2y - e1y * e2x;
			const float l2 = nx * nx + ny * ny + nz * nz, il = l2 > 0 ? 1.0f / sqrtf( l2 ) : 0.f;
			float4* sp = a.shade + a.triBase + prim;
			*sp = make_float4( nx * il, ny * il, nz * il, li ? __uint_as_float( words[k] ) : sp->w );
		}
		lo[0] = fminf( lo[0], fminf( p0.x, fminf( p1.x, p2.x ) ) ), lo[1] = fminf( lo[1], fminf( p0.y, fminf( p1.y, p2.y ) ) );
		lo[2] = fminf( lo[2], fminf( p0.z, fminf( p1.z, p2.z ) ) );
		hi[0] = fmaxf( hi[0], fmaxf( p0.x, fmaxf( p1.x, p2.x ) ) ), hi[1] = fmaxf( hi[1], fmaxf( p0.y, fmaxf( p1.y, p2.y ) ) );
		hi[2] = fmaxf( hi[2], fmaxf( p0.z, fmaxf( p1.z, p2.z ) ) );
	}
}

// the boxes of the eight slots are known (lane s holds slot s; empty = inverted box): node box by an 8-lane shuffle reduction, the
// node's frame, every lane's planes against it; lane 0 writes the frame and the float box for the parent.  All 8 lanes of the group.
__device__ __forceinline__ void finishNode( const RefitArgs& a, const uint32_t n, const uint32_t s, const uint32_t mask, const uint32_t imask, const uint32_t meta,
	const float lo[3], const float hi[3], const bool valid = true )
{
	float nlo[3], nhi[3];
	for (int c = 0; c < 3; c++)
	{
		float l = lo[c], h = hi[c];
		for (int d = 1; d < 8; d <<= 1) l = fminf( l, __shfl_xor_sync( mask, l, d ) ), h = fmaxf( h, __shfl_xor_sync( mask, h, d ) );
		nlo[c] = l, nhi[c] = h;
	}
	if (!valid) return;
	uint4* np = a.nodes + (size_t)(a.nodeBase + n) * 5;
	uint32_t e[3], qlo[3] = { 0, 0, 0 }, qhi[3] = { 0, 0, 0 };
	float fp[3];
	for (int c = 0; c < 3; c++) e[c] = frameOf( nlo[c], nhi[c], fp[c] );
	if (meta) for (int c = 0; c < 3; c++) quantise( fp[c], e[c], lo[c], hi[c], qlo[c], qhi[c] );
	uint8_t* q = reinterpret_cast<uint8_t*>( np + 2 ); // qlox[8] qloy[8] qloz[8] qhix[8] qhiy[8] qhiz[8]
	q[s] = (uint8_t)qlo[0], q[8 + s] = (uint8_t)qlo[1], q[16 + s] = (uint8_t)qlo[2];
	q[24 + s] = (uint8_t)qhi[0], q[32 + s] = (uint8_t)qhi[1], q[40 + s] = (uint8_t)qhi[2];
	if (s == 0)
	{
		np[0] = make_uint4( __float_as_uint( fp[0] ), __float_as_uint( fp[1] ), __float_as_uint( fp[2] ),
			e[0] | (e[1] << 8) | (e[2] << 16) | (imask << 24) );
		a.nodeLo[n] = make_float4( nlo[0], nlo[1], nlo[2], 0 ), a.nodeHi[n] = make_float4( nhi[0], nhi[1], nhi[2], 0 );
	}
}

// one child slot of one node, level by level (the children's boxes come from an earlier launch or an earlier barrier); all 8 lanes of a
// node's group call this together (valid or not), groups are 8-aligned in the warp
__device__ __forceinline__ void refitSlot( const RefitArgs& a, const uint32_t node, const uint32_t s, const bool valid, const uint32_t levelBegin )
{
	const uint32_t n = valid ? node : levelBegin;
	const uint4* np = a.nodes + (size_t)(a.nodeBase + n) * 5;
	const uint32_t imask = np[0].w >> 24;
	const uint4 n1 = np[1];
	const uint32_t meta = ((s < 4 ? n1.z : n1.w) >> (8 * (s & 3))) & 0xffu;
	float lo[3] = { 3.0e38f, 3.0e38f, 3.0e38f }, hi[3] = { -3.0e38f, -3.0e38f, -3.0e38f };
	if (meta)
	{
		if (imask & (1u << s))
		{
			// absolute child index -> index relative to this mesh's scratch
			const uint32_t child = n1.x + __popc( imask & ((1u << s) - 1u) ) - a.nodeBase;
			const float4 cl = a.nodeLo[child], ch = a.nodeHi[child];
			lo[0] = cl.x, lo[1] = cl.y, lo[2] = cl.z, hi[0] = ch.x, hi[1] = ch.y, hi[2] = ch.z;
		}
		else leafSlotBox( a, n1, meta, lo, hi );
	}
	finishNode( a, n, s, 0xffffffffu, imask, meta, lo, hi, valid );
}

__global__ void refitLevelKernel( const RefitArgs a, const uint32_t levelBegin, const uint32_t levelEnd )
{
	const uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x;
	const uint32_t node = levelBegin + (gid >> 3);
	refitSlot( a, node, gid & 7, node < levelEnd, levelBegin ); // whole 8-lane groups are valid or not, so the shuffles stay converged per group
}

// the small top levels in ONE launch: a single CTA walks them bottom-up with a block barrier between levels
struct TopLevels { uint32_t start[17]; int count; }; // start[i]..start[i+1] = level i (root level first)
__global__ void __launch_bounds__( 1024 ) refitTopKernel( const RefitArgs a, const TopLevels t )
{
	for (int level = t.count - 1; level >= 0; level--)
	{
		const uint32_t begin = t.start[level], end = t.start[level + 1], slots = (end - begin) * 8;
		for (uint32_t i = threadIdx.x; i < ((slots + 1023u) & ~1023u); i += 1024)
		{
			const uint32_t node = begin + (i >> 3);
			refitSlot( a, node, i & 7, node < end, begin );
		}
		__threadfence_block();
		__syncthreads();
	}
}

// leaf-order copy of (primitive index, the word the shading record carries beside the normal) for leafSlotBox
__global__ void refitLeafInfoKernel( const float4* __restrict__ tris, const float4* __restrict__ shade, const uint32_t triBase, const uint32_t triCount, uint2* __restrict__ out )
{
	const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
	if (j >= triCount) return;
	const uint32_t prim = __float_as_uint( tris[(size_t)(triBase + j) * 3].w );
	out[j] = make_uint2( prim, shade ? __float_as_uint( shade[triBase + prim].w ) : 0u );
}

// parentInfo[n] = parent of node n (relative to the mesh's root; NO_PARENT for the root) | number of inner children of that parent << 28;
// stashOf[n] = a dense number for every node WITH inner children (its place in the walk's slot-box stash); *stashCount = how many
constexpr uint32_t NO_PARENT = 0x0fffffffu;
__global__ void refitParentsKernel( const uint4* __restrict__ nodes, const uint32_t nodeBase, const uint32_t nodeCount, uint32_t* __restrict__ parentInfo, uint32_t* __restrict__ arrived,
	uint32_t* __restrict__ stashOf, uint32_t* stashCount )
{
	const uint32_t n = blockIdx.x * blockDim.x + threadIdx.x;
	if (n >= nodeCount) return;
	if (n == 0) parentInfo[0] = NO_PARENT;
	arrived[n] = 0;
	const uint4* np = nodes + (size_t)(nodeBase + n) * 5;
	const uint32_t imask = np[0].w >> 24, inner = __popc( imask ), first = np[1].x - nodeBase;
	for (uint32_t k = 0; k < inner; k++) parentInfo[first + k] = n | (inner << 28);
	stashOf[n] = inner ? atomicAdd( stashCount, 1u ) : 0xffffffffu;
}

// The whole tree in one launch: see the file comment.  Eight lanes per node; the four groups of a warp go their own ways (group masks).
// Every group first does the triangles of its node's leaf slots -- that is the streaming part, and nothing in it waits for anybody.
// A node without inner children is finished with that.  A node WITH inner children parks its leaf-slot boxes in the stash and counts as
// one more arrival at itself; whoever arrives last at a node -- its own group or the group that brings the last child -- finishes it
// (child boxes from the scratch boxes, leaf-slot boxes from the stash) and goes on to the parent.
#ifndef REFIT_MIN_BLOCKS
#define REFIT_MIN_BLOCKS 4
#endif
__global__ void __launch_bounds__( 256, REFIT_MIN_BLOCKS ) refitWalkKernel( const RefitArgs a, const uint32_t nodeCount, const uint32_t* __restrict__ parentInfo, uint32_t* arrived,
	const uint32_t* __restrict__ stashOf, float4* stash )
{
	const uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x, lane = threadIdx.x & 31, s = lane & 7;
	const uint32_t mask = 0xffu << (lane & 24);
	if ((gid >> 3) >= nodeCount) return; // (whole groups)
	// deepest nodes first (nodes are stored breadth-first): their way up is the longest and runs beside the triangle work of the shallower ones
	uint32_t node = nodeCount - 1 - (gid >> 3);
	const uint4* np = a.nodes + (size_t)(a.nodeBase + node) * 5;
	uint32_t imask = np[0].w >> 24;
	uint4 n1 = np[1];
	uint32_t info = __ldg( parentInfo + node );
	uint32_t meta = ((s < 4 ? n1.z : n1.w) >> (8 * (s & 3))) & 0xffu;
	float lo[3] = { 3.0e38f, 3.0e38f, 3.0e38f }, hi[3] = { -3.0e38f, -3.0e38f, -3.0e38f };
	if (meta && !(imask & (1u << s))) leafSlotBox( a, n1, meta, lo, hi );
	bool own = true; // the slot boxes in lo / hi are this group's own work (not yet: the children's)
	if (imask)
	{
		float4* st = stash + (size_t)__ldg( stashOf + node ) * 16;
		st[s] = make_float4( lo[0], lo[1], lo[2], 0 ), st[8 + s] = make_float4( hi[0], hi[1], hi[2], 0 );
		__syncwarp( mask );
		uint32_t last = 0;
		if (s == 0)
		{
			__threadfence(); // (cumulative: the seven other lanes' stash stores are ordered before it by the barrier)
			last = atomicAdd( arrived + node, 1u ) == __popc( imask ) ? 1u : 0u;
			if (last) __threadfence(), arrived[node] = 0; // (the fence before the reset: it then has no store of ours to wait for)
		}
		if (!__shfl_sync( mask, last, lane & 24 )) return;
	}
	while (true)
	{
		// (a) finish `node`: every child has delivered.  Inner slots take the child's box; leaf slots have theirs (own) or fetch it from the stash
		if (meta)
		{
			if (imask & (1u << s))
			{
				const uint32_t child = n1.x + __popc( imask & ((1u << s) - 1u) ) - a.nodeBase; // absolute child index -> relative to this mesh's scratch
				const float4 cl = __ldcg( a.nodeLo + child ), ch = __ldcg( a.nodeHi + child );
				lo[0] = cl.x, lo[1] = cl.y, lo[2] = cl.z, hi[0] = ch.x, hi[1] = ch.y, hi[2] = ch.z;
			}
			else if (!own)
			{
				const float4* st = stash + (size_t)__ldg( stashOf + node ) * 16;
				const float4 cl = __ldcg( st + s ), ch = __ldcg( st + 8 + s );
				lo[0] = cl.x, lo[1] = cl.y, lo[2] = cl.z, hi[0] = ch.x, hi[1] = ch.y, hi[2] = ch.z;
			}
		}
		// the way up is a chain of dependent round trips to L2; what does not depend on this node's result is asked for before it:
		// the parent's topology words and its own parent link (nobody writes them before the last arrival -- and that may be us)
		const uint32_t parent = info & NO_PARENT;
		uint32_t pn0w = 0, pinfo = 0;
		uint4 pn1 = make_uint4( 0, 0, 0, 0 );
		if (parent != NO_PARENT)
		{
			const uint4* pp = a.nodes + (size_t)(a.nodeBase + parent) * 5;
			pn0w = __ldcg( &pp[0].w ), pn1 = __ldcg( pp + 1 ), pinfo = __ldg( parentInfo + parent );
		}
		finishNode( a, node, s, mask, imask, meta, lo, hi );
		if (parent == NO_PARENT) return;
		// (b) deliver: the node's box (lane 0 wrote it) must be visible before the arrival is; the planes and the frame are read by nobody in this launch
		uint32_t last = 0;
		if (s == 0)
		{
			__threadfence();
			last = atomicAdd( arrived + parent, 1u ) == (info >> 28) ? 1u : 0u; // (inner children + the parent's own group)
			if (last) __threadfence(), arrived[parent] = 0; // the siblings' boxes are read after this; counter ready for the next refit (nobody else touches it any more)
		}
		if (!__shfl_sync( mask, last, lane & 24 )) return;
		node = parent, imask = pn0w >> 24, n1 = pn1, info = pinfo, own = false;
		meta = ((s < 4 ? n1.z : n1.w) >> (8 * (s & 3))) & 0xffu;
		lo[0] = lo[1] = lo[2] = 3.0e38f, hi[0] = hi[1] = hi[2] = -3.0e38f;
	}
}

} // namespace

cudaError_t LaunchRefitParents( const uint4* nodes, uint32_t nodeBase, uint32_t nodeCount, uint32_t* parentInfo, uint32_t* arrived, uint32_t* stashOf, uint32_t* stashCount, cudaStream_t s )
{
	if (!nodeCount) return cudaSuccess;
	cudaError_t e = cudaMemsetAsync( stashCount, 0, 4, s );
	if (e != cudaSuccess) return e;
	refitParentsKernel<<<(nodeCount + 255) / 256, 256, 0, s>>>( nodes, nodeBase, nodeCount, parentInfo, arrived, stashOf, stashCount );
	return cudaGetLastError();
}

cudaError_t LaunchRefitLeafInfo( const float4* tris, const float4* shade, uint32_t triBase, uint32_t triCount, uint2* out, cudaStream_t s )
{
	if (!triCount) return cudaSuccess;
	refitLeafInfoKernel<<<(triCount + 255) / 256, 256, 0, s>>>( tris, shade, triBase, triCount, out );
	return cudaGetLastError();
}

cudaError_t LaunchRefitWalk( const RefitArgs& a, uint32_t nodeCount, const uint32_t* parentInfo, uint32_t* arrived, const uint32_t* stashOf, float4* stash, cudaStream_t s )
{
	if (!nodeCount) return cudaSuccess;
	refitWalkKernel<<<(nodeCount * 8 + 255) / 256, 256, 0, s>>>( a, nodeCount, parentInfo, arrived, stashOf, stash );
	return cudaGetLastError();
}

// levels [0, nTop) of `levelStart` (root first) in one launch; nTop <= 16
cudaError_t LaunchRefitTop( const RefitArgs& a, const uint32_t* levelStart, int nTop, cudaStream_t s )
{
	if (nTop <= 0) return cudaSuccess;
	TopLevels t;
	t.count = nTop;
	for (int i = 0; i <= nTop; i++) t.start[i] = levelStart[i];
	refitTopKernel<<<1, 1024, 0, s>>>( a, t );
	return cudaGetLastError();
}

cudaError_t LaunchRefitLevel( const RefitArgs& a, uint32_t levelBegin, uint32_t levelEnd, cudaStream_t s )
{
	const uint32_t threads = (levelEnd - levelBegin) * 8;
	if (!threads) return cudaSuccess;
	refitLevelKernel<<<(threads + 255) / 256, 256, 0, s>>>( a, levelBegin, levelEnd );
	return cudaGetLastError();
}

} // namespace b200
