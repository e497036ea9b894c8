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
		// frexp( ext / 254.75 ): the exponent field of the double (no denormals here: ext comes from floats), + 1 for the [0.5, 1) mantissa
		const int fe = ((__double2hiint( ext / 254.75 ) >> 20) & 0x7ff) - 1022;
		ex = fe;
		if (pow2( ex - 1 ) * 254.75 >= ext) ex--;
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

// one child slot of one node; all 8 lanes of a node's group call this together (valid or not), groups are 8-aligned in the warp
template <bool COHERENT /* child boxes come from groups of the same launch: read them from L2 */>
__device__ __forceinline__ void refitSlot( const RefitArgs& a, const uint32_t node, const uint32_t s, const bool valid, const uint32_t levelBegin, const uint32_t mask = 0xffffffffu )
{
	const uint32_t n = valid ? node : levelBegin;
	uint4* np = a.nodes + (size_t)(a.nodeBase + n) * 5;
	const uint4 n0 = np[0], n1 = np[1];
	const uint32_t imask = n0.w >> 24;
	const uint32_t meta = ((s < 4 ? n1.z : n1.w) >> (8 * (s & 3))) & 0xffu;
	float lo[3] = { 3.0e38f, 3.0e38f, 3.0e38f }, hi[3] = { -3.0e38f, -3.0e38f, -3.0e38f };
	if (meta)
	{
		if (imask & (1u << s))
		{
			// absolute child index -> index relative to this mesh's scratch
			const uint32_t child = n1.x + __popc( imask & ((1u << s) - 1u) ) - a.nodeBase;
			const float4 cl = COHERENT ? __ldcg( a.nodeLo + child ) : a.nodeLo[child], ch = COHERENT ? __ldcg( a.nodeHi + child ) : a.nodeHi[child];
			lo[0] = cl.x, lo[1] = cl.y, lo[2] = cl.z, hi[0] = ch.x, hi[1] = ch.y, hi[2] = ch.z;
		}
		else
		{
			// leaf slot: this lane owns the slot's <= 3 triangles.  Gather their new vertices ONCE (original primitive order),
			// rewrite the leaf-order records {v0|prim, e1, e2} and the shading normal, and bound them.
			const uint32_t cnt = __popc( meta >> 5 ), first = n1.y + (meta & 31u);
			for (uint32_t k = 0; k < cnt; k++)
			{
				float4* tp = a.tris + (size_t)(first + k) * 3;
				const uint32_t prim = __float_as_uint( tp[0].w );
				const float4* vp = a.verts + (size_t)prim * 3;
				const float4 p0 = __ldg( vp ), p1 = __ldg( vp + 1 ), p2 = __ldg( vp + 2 );
				const float e1x = p1.x - p0.x, e1y = p1.y - p0.y, e1z = p1.z - p0.z, e2x = p2.x - p0.x, e2y = p2.y - p0.y, e2z = p2.z - p0.z;
				tp[0] = make_float4( p0.x, p0.y, p0.z, __uint_as_float( prim ) );
				tp[1] = make_float4( e1x, e1y, e1z, 0.f );
				tp[2] = make_float4( e2x, e2y, e2z, 0.f );
				if (a.shade)
				{
					float nx = e1y * e2z - e1z * e2y, ny = e1z * e2x - e1x * e2z, nz = e1x * e2y - e1y * e2x;
					const float l2 = nx * nx + ny * ny + nz * nz, il = l2 > 0 ? 1.0f / sqrtf( l2 ) : 0.f;
					float4* sp = a.shade + a.triBase + prim;
					*sp = make_float4( nx * il, ny * il, nz * il, sp->w );
				}
				lo[0] = fminf( lo[0], fminf( p0.x, fminf( p1.x, p2.x ) ) ), lo[1] = fminf( lo[1], fminf( p0.y, fminf( p1.y, p2.y ) ) );
				lo[2] = fminf( lo[2], fminf( p0.z, fminf( p1.z, p2.z ) ) );
				hi[0] = fmaxf( hi[0], fmaxf( p0.x, fmaxf( p1.x, p2.x ) ) ), hi[1] = fmaxf( hi[1], fmaxf( p0.y, fmaxf( p1.y, p2.y ) ) );
				hi[2] = fmaxf( hi[2], fmaxf( p0.z, fmaxf( p1.z, p2.z ) ) );
			}
		}
	}
	float nlo[3], nhi[3];
	for (int c = 0; c < 3; c++)
	{
		float l = lo[c], h = hi[c];
		for (int d = 1; d < 8; d <<= 1) l = fminf( l, __shfl_xor_sync( mask, l, d ) ), h = fmaxf( h, __shfl_xor_sync( mask, h, d ) );
		nlo[c] = l, nhi[c] = h;
	}
	if (!valid) return;
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

__global__ void refitLevelKernel( const RefitArgs a, const uint32_t levelBegin, const uint32_t levelEnd )
{
	const uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x;
	const uint32_t node = levelBegin + (gid >> 3);
	refitSlot<false>( a, node, gid & 7, node < levelEnd, levelBegin ); // whole 8-lane groups are valid or not, so the shuffles stay converged per group
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
			refitSlot<false>( a, node, i & 7, node < end, begin );
		}
		__threadfence_block();
		__syncthreads();
	}
}

// parentInfo[n] = parent of node n (relative to the mesh's root; NO_PARENT for the root) | number of inner children of that parent << 28
constexpr uint32_t NO_PARENT = 0x0fffffffu;
__global__ void refitParentsKernel( const uint4* __restrict__ nodes, const uint32_t nodeBase, const uint32_t nodeCount, uint32_t* __restrict__ parentInfo, uint32_t* __restrict__ arrived )
{
	const uint32_t n = blockIdx.x * blockDim.x + threadIdx.x;
	if (n >= nodeCount) return;
	if (n == 0) parentInfo[0] = NO_PARENT;
	arrived[n] = 0;
	const uint4* np = nodes + (size_t)(nodeBase + n) * 5;
	const uint32_t imask = np[0].w >> 24, inner = __popc( imask ), first = np[1].x - nodeBase;
	for (uint32_t k = 0; k < inner; k++) parentInfo[first + k] = n | (inner << 28);
}

// the whole tree in one launch: see the file comment.  Eight lanes per node; the four groups of a warp climb independently (group masks).
__global__ void __launch_bounds__( 256 ) refitWalkKernel( const RefitArgs a, const uint32_t nodeCount, const uint32_t* __restrict__ parentInfo, uint32_t* arrived )
{
	const uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x, lane = threadIdx.x & 31, s = lane & 7;
	const uint32_t mask = 0xffu << (lane & 24);
	uint32_t node = gid >> 3;
	if (node >= nodeCount) return; // (whole groups)
	if (a.nodes[(size_t)(a.nodeBase + node) * 5].w >> 24) return; // has inner children: the group that brings the last of them continues here
	while (true)
	{
		refitSlot<true>( a, node, s, true, 0, mask );
		const uint32_t info = __ldg( parentInfo + node ), parent = info & NO_PARENT;
		if (parent == NO_PARENT) return;
		// the node's box and planes must be visible before the arrival is: every lane fences its own stores, then one lane counts
		__threadfence();
		__syncwarp( mask );
		uint32_t last = 0;
		if (s == 0)
		{
			last = atomicAdd( arrived + parent, 1u ) + 1 == (info >> 28) ? 1u : 0u;
			if (last) arrived[parent] = 0; // ready for the next refit (nobody else touches this counter any more)
		}
		last = __shfl_sync( mask, last, lane & 24 );
		if (!last) return;
		__threadfence();
		node = parent;
	}
}

} // namespace

cudaError_t LaunchRefitParents( const uint4* nodes, uint32_t nodeBase, uint32_t nodeCount, uint32_t* parentInfo, uint32_t* arrived, cudaStream_t s )
{
	if (!nodeCount) return cudaSuccess;
	refitParentsKernel<<<(nodeCount + 255) / 256, 256, 0, s>>>( nodes, nodeBase, nodeCount, parentInfo, arrived );
	return cudaGetLastError();
}

cudaError_t LaunchRefitWalk( const RefitArgs& a, uint32_t nodeCount, const uint32_t* parentInfo, uint32_t* arrived, cudaStream_t s )
{
	if (!nodeCount) return cudaSuccess;
	refitWalkKernel<<<(nodeCount * 8 + 255) / 256, 256, 0, s>>>( a, nodeCount, parentInfo, arrived );
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
