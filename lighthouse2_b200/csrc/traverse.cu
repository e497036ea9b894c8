// traverse.cu -- the extend (closest-hit) and shadow (any-hit) waves of RenderCore_B200.
//
// These two kernels replace, on sm_100a, what the reference cores delegate to closed libraries:
//   extend : optixTrace closest-hit (lib/rendercore_optix7/optix/.optix.cu:125-135,177-187),
//            rtpQueryExecute CLOSEST (lib/RenderCore_OptixPrime_B/rendercore.cpp:524-528)
//   shadow : optixTrace TERMINATE_ON_FIRST_HIT (.optix.cu:137-149), rtpQueryExecute ANY (:556-565)
// and, for results, what the CPU core computes in TopLevelBVH::Traverse / BVH2::Traverse /
// Ray::IntersectTriangle (lib/RenderCore_Bart/bvh.cpp:594,258; common.h:19).
//
// Design (B200-first; no tensor cores: traversal is not a contraction):
//  * persistent warps; lanes fetch rays with one warp-aggregated atomicAdd, refilling idle lanes when
//    too few are active (ray-stream compaction by vote + prefix popcount)
//  * rays are two coalesced float4 planes; hits are one float4
//  * 8-wide BVH with 8-bit quantised child boxes (wide_bvh.h); five LDG.128 per node; the eight slab
//    tests decode planes with PRMT (+ HADD2.F32) and one FFMA each; octant-ordered child visiting from a hit mask
//  * leaf triangles go through a per-warp ring in shared memory and are tested 32 at a time, one per lane
//  * traversal stack: (base,mask) groups, first STACK_SMEM entries per lane in shared memory, rest local
//  * TLAS -> BLAS: an instance leaf transforms the ray (reference arithmetic in ISECT_REFERENCE mode)
//    and traversal continues in the same loop; leaving the instance restores the world ray.  Scenes of
//    one instance run kernel instances (SINGLE) that take the ray into the instance when it is fetched
//  * a pass holds what every ray needs every step and nothing else: results are written and rays fetched
//    between rounds of passes, the warp votes once per pass and reads the vote one pass later
//
// Compiled with -fmad=false: every float op below rounds separately unless written as __fmaf_rn, so the
// ISECT_REFERENCE triangle test and instance entry are bit-identical to the reference's scalar C++.
#include "kernels.h"
#include "wide_bvh.h"
#include <cuda_fp16.h>

#ifndef B200_HALF_DECODE
#define B200_HALF_DECODE 1 // node planes: two bytes per PRMT + HADD2.F32 (1) or one PRMT per byte (0); see qpair / qbyte.  72 / 48 decode
                            // instructions per node, and a 60-instruction mix of the two, all trace within 1 % of each other: the kernel is bound by
                            // the dependent chain of a pass at 6 warps per scheduler, not by instruction issue
#endif

namespace b200 {

namespace {

constexpr int BLOCK_THREADS = 128;
#ifndef B200_STACK_SMEM
#define B200_STACK_SMEM 8
#endif
constexpr int STACK_SMEM = B200_STACK_SMEM; // per-lane stack entries kept in shared memory
constexpr int STACK_LOCAL = TRAVERSAL_STACK_ENTRIES - STACK_SMEM; // overflow entries in local memory; the host checks tree depths against the sum (wide_bvh.h)
constexpr float REF_EPSILON = 0.0001f; // lib/RenderSystem/common_settings.h:45
constexpr uint32_t NO_INSTANCE = 0xffffffffu;

__device__ __forceinline__ float max3f( float a, float b, float c )
{
	float r;
	asm( "max.f32 %0, %1, %2, %3;" : "=f"( r ) : "f"( a ), "f"( b ), "f"( c ) ); // FMNMX3 (sm_100+)
	return r;
}
__device__ __forceinline__ float min3f( float a, float b, float c )
{
	float r;
	asm( "min.f32 %0, %1, %2, %3;" : "=f"( r ) : "f"( a ), "f"( b ), "f"( c ) );
	return r;
}
__device__ __forceinline__ uint32_t lanemask_lt()
{
	uint32_t m;
	asm( "mov.u32 %0, %%lanemask_lt;" : "=r"( m ) );
	return m;
}
__device__ __forceinline__ uint32_t sign_extend_s8x4( uint32_t x )
{
	// each byte: 0x80 set -> 0xff, else 0x00
	uint32_t r;
	asm( "prmt.b32 %0, %1, 0, 0xba98;" : "=r"( r ) : "r"( x ) ); // replicate-sign mode selectors 8..b
	return r;
}
// byte J of q placed in byte 1 of K = 0x4B000000: the float 2^23 + 256*b (exact), decoded by one FFMA.  K arrives as
// a kernel argument so that ptxas keeps it in a register and encodes the SELECTOR as the immediate (PRMT has one
// immediate slot; with a literal K the four selectors were re-materialised with a MOV in front of every PRMT).
template <int J> __device__ __forceinline__ float qbyte( uint32_t q, uint32_t K )
{
	uint32_t r;
	if (J == 0) asm( "prmt.b32 %0, %1, %2, 0x7604;" : "=r"( r ) : "r"( q ), "r"( K ) );
	else if (J == 1) asm( "prmt.b32 %0, %1, %2, 0x7614;" : "=r"( r ) : "r"( q ), "r"( K ) );
	else if (J == 2) asm( "prmt.b32 %0, %1, %2, 0x7624;" : "=r"( r ) : "r"( q ), "r"( K ) );
	else asm( "prmt.b32 %0, %1, %2, 0x7634;" : "=r"( r ) : "r"( q ), "r"( K ) );
	return __uint_as_float( r );
}

// Two bytes per PRMT: bytes 2P and 2P+1 of q become the half2 (1024 + b0, 1024 + b1) -- 0x64 from K = 0x64646464 is the high
// byte of each half -- and two HADD2.F32 turn them into floats.  The conversions run on the FMA pipe, the PRMT on the ALU pipe,
// which is the one the node test saturates (145 ALU-pipe vs 80 FMA-pipe instructions per node with one PRMT per byte; both
// pipes take one warp instruction every two cycles).
template <int P> __device__ __forceinline__ void qpair( const uint32_t q, const uint32_t K, float& f0, float& f1 )
{
	uint32_t r;
	if (P == 0) asm( "prmt.b32 %0, %1, %2, 0x4140;" : "=r"( r ) : "r"( q ), "r"( K ) );
	else asm( "prmt.b32 %0, %1, %2, 0x4342;" : "=r"( r ) : "r"( q ), "r"( K ) );
	const __half2 h = *reinterpret_cast<const __half2*>( &r );
	f0 = __low2float( h ), f1 = __high2float( h );
}

// m with its k highest set bits cleared (k < popc(m)).  A branch-free halving search and fns.b32 (45 SASS instructions) were
// looked at: both have a longer dependent chain than this loop at the k <= 5 that is typical
__device__ __forceinline__ uint32_t dropHighest( uint32_t m, uint32_t k )
{
	for (; k; k--) m &= ~(0x80000000u >> __clz( m ));
	return m;
}

struct RayState {
	float ox, oy, oz, dx, dy, dz;   // current-space ray (world in the TLAS, object inside an instance)
	float ix, iy, iz;               // 1/d with zero components replaced by +-1e-18 (slab test only)
	float tmin, t;                  // t = current closest hit / tmax
	uint32_t octinv4;               // (7 - octant) replicated to 4 bytes
};

__device__ __forceinline__ void setupSlab( RayState& r )
{
	const float tiny = 1e-18f;
	float dx = fabsf( r.dx ) > tiny ? r.dx : copysignf( tiny, r.dx );
	float dy = fabsf( r.dy ) > tiny ? r.dy : copysignf( tiny, r.dy );
	float dz = fabsf( r.dz ) > tiny ? r.dz : copysignf( tiny, r.dz );
	r.ix = __frcp_rn( dx ), r.iy = __frcp_rn( dy ), r.iz = __frcp_rn( dz );
	uint32_t oct = (r.dx < 0 ? 4u : 0u) | (r.dy < 0 ? 2u : 0u) | (r.dz < 0 ? 1u : 0u);
	r.octinv4 = (7u - oct) * 0x01010101u;
}

// Instance entry.  ISECT_REFERENCE: exactly TopLevelBVH::Traverse's leaf code (bvh.cpp:607-610):
//   O' = inv.TransformPoint(O)  (common_types.h:623-635, incl. the w != 1 divide)
//   D' = normalize(inv.TransformVector(D))  (common_types.h:616-621; normalize = v * (1.0f/sqrtf(dot)), :87,363)
// so t inside the instance is measured along the re-normalised object-space direction, as in the reference.
// ISECT_WATERTIGHT: affine transform, direction NOT renormalised (t stays the world-space parameter).
template <int ISECT>
__device__ __forceinline__ void enterInstance( RayState& r, const float4* __restrict__ rec )
{
	const float4 r0 = __ldg( rec + 0 ), r1 = __ldg( rec + 1 ), r2 = __ldg( rec + 2 );
	const float ox = r.ox, oy = r.oy, oz = r.oz, dx = r.dx, dy = r.dy, dz = r.dz;
	if (ISECT == ISECT_REFERENCE)
	{
		const float4 r3 = __ldg( rec + 3 );
		float px = r0.x * ox + r0.y * oy + r0.z * oz + r0.w;
		float py = r1.x * ox + r1.y * oy + r1.z * oz + r1.w;
		float pz = r2.x * ox + r2.y * oy + r2.z * oz + r2.w;
		float w = r3.x * ox + r3.y * oy + r3.z * oz + r3.w;
		if (w != 1) { float iw = 1.f / w; px = px * iw, py = py * iw, pz = pz * iw; }
		float vx = r0.x * dx + r0.y * dy + r0.z * dz;
		float vy = r1.x * dx + r1.y * dy + r1.z * dz;
		float vz = r2.x * dx + r2.y * dy + r2.z * dz;
		float invLen = 1.0f / sqrtf( vx * vx + vy * vy + vz * vz );
		r.ox = px, r.oy = py, r.oz = pz;
		r.dx = vx * invLen, r.dy = vy * invLen, r.dz = vz * invLen;
	}
	else
	{
		r.ox = __fmaf_rn( r0.x, ox, __fmaf_rn( r0.y, oy, __fmaf_rn( r0.z, oz, r0.w ) ) );
		r.oy = __fmaf_rn( r1.x, ox, __fmaf_rn( r1.y, oy, __fmaf_rn( r1.z, oz, r1.w ) ) );
		r.oz = __fmaf_rn( r2.x, ox, __fmaf_rn( r2.y, oy, __fmaf_rn( r2.z, oz, r2.w ) ) );
		r.dx = __fmaf_rn( r0.x, dx, __fmaf_rn( r0.y, dy, r0.z * dz ) );
		r.dy = __fmaf_rn( r1.x, dx, __fmaf_rn( r1.y, dy, r1.z * dz ) );
		r.dz = __fmaf_rn( r2.x, dx, __fmaf_rn( r2.y, dy, r2.z * dz ) );
	}
	setupSlab( r );
}

// Triangle test.  ISECT_REFERENCE restates Ray::IntersectTriangle (lib/RenderCore_Bart/common.h:19-50)
// operation by operation (e1/e2 are the same floats the reference computes at :25-26; f = 1.0/a in double
// rounded to float equals the IEEE float quotient; bounds (EPSILON, 1/EPSILON) at :43), plus the wave's
// own (tmin, t) interval.  O = (ox,oy,oz,tmin), D = (dx,dy,dz,t).  Returns true and writes tt,u,v on acceptance
// (strict found_t < t, bvh.cpp:268).
// Exact ties (two triangles with bit-identical t: a ray through a shared edge) go to the triangle with the smaller leaf-order index
// `tri`, whatever order the warp happens to test them in: `bestTri` is the index of the owner's current hit.  (The reference keeps
// the triangle ITS tree visits first, bvh.cpp:268 -- an order no other tree reproduces, so ties are a documented exception class of
// the parity checker either way; a rule that does not depend on which rays share a warp makes the wave's output reproducible from
// run to run, which the bit-identical multi-GPU frame check needs.)
template <int ISECT>
__device__ __forceinline__ bool intersectTriangle( const float4 O, const float4 D, const float4 v0, const float4 e1, const float4 e2, const uint32_t tri,
	const uint32_t* bestTri, float& tOut, float& uOut, float& vOut )
{
	if (ISECT == ISECT_REFERENCE)
	{
		const float hx = D.y * e2.z - D.z * e2.y, hy = D.z * e2.x - D.x * e2.z, hz = D.x * e2.y - D.y * e2.x;
		const float a = e1.x * hx + e1.y * hy + e1.z * hz;
		if (a > -REF_EPSILON && a < REF_EPSILON) return false;
		const float f = 1.0f / a;
		const float sx = O.x - v0.x, sy = O.y - v0.y, sz = O.z - v0.z;
		const float u = f * (sx * hx + sy * hy + sz * hz);
		if (u < 0.0f || u > 1.0f) return false;
		const float qx = sy * e1.z - sz * e1.y, qy = sz * e1.x - sx * e1.z, qz = sx * e1.y - sy * e1.x;
		const float v = f * (D.x * qx + D.y * qy + D.z * qz);
		if (v < 0.0f || u + v > 1.0f) return false;
		const float tt = f * (e2.x * qx + e2.y * qy + e2.z * qz);
		if (tt > REF_EPSILON && tt < 1 / REF_EPSILON && tt > O.w && (tt < D.w || (tt == D.w && tri < *bestTri))) { tOut = tt, uOut = u, vOut = v; return true; }
		return false;
	}
	else
	{
		// fused Moeller-Trumbore without absolute epsilons (scale invariant); edges shared by two
		// triangles see the same rounded products, inclusive on both sides
		const float hx = __fmaf_rn( D.y, e2.z, -(D.z * e2.y) ), hy = __fmaf_rn( D.z, e2.x, -(D.x * e2.z) ),
					hz = __fmaf_rn( D.x, e2.y, -(D.y * e2.x) );
		const float a = __fmaf_rn( e1.x, hx, __fmaf_rn( e1.y, hy, e1.z * hz ) );
		if (a == 0.0f) return false;
		const float f = __frcp_rn( a );
		const float sx = O.x - v0.x, sy = O.y - v0.y, sz = O.z - v0.z;
		const float u = f * __fmaf_rn( sx, hx, __fmaf_rn( sy, hy, sz * hz ) );
		if (u < 0.0f || u > 1.0f) return false;
		const float qx = __fmaf_rn( sy, e1.z, -(sz * e1.y) ), qy = __fmaf_rn( sz, e1.x, -(sx * e1.z) ),
					qz = __fmaf_rn( sx, e1.y, -(sy * e1.x) );
		const float v = f * __fmaf_rn( D.x, qx, __fmaf_rn( D.y, qy, D.z * qz ) );
		if (v < 0.0f || u + v > 1.0f) return false;
		const float tt = f * __fmaf_rn( e2.x, qx, __fmaf_rn( e2.y, qy, e2.z * qz ) );
		if (tt > O.w && (tt < D.w || (tt == D.w && tri < *bestTri)) && tt > 0.0f) { tOut = tt, uOut = u, vOut = v; return true; }
		return false;
	}
}


// One wide-node visit: decode the 8 quantised child boxes and return the octant-ordered hit mask
// (bits 31..24: inner children in visiting order, bits 23..0: leaf primitives relative to primBase).
template <int J> __device__ __forceinline__ void testChild( const RayState& r, const float tcull, const uint32_t K, const uint32_t nearx, const uint32_t neary,
	const uint32_t nearz, const uint32_t farx, const uint32_t fary, const uint32_t farz, const float ax, const float ay, const float az,
	const float bx, const float by, const float bz, const float axf, const float ayf, const float azf, const float bxf, const float byf,
	const float bzf, const uint32_t childBits4, const uint32_t bitIndex4, uint32_t& hitmask )
{
	const float tnx = __fmaf_rn( qbyte<J>( nearx, K ), ax, bx ), tny = __fmaf_rn( qbyte<J>( neary, K ), ay, by ),
				tnz = __fmaf_rn( qbyte<J>( nearz, K ), az, bz );
	const float tfx = __fmaf_rn( qbyte<J>( farx, K ), axf, bxf ), tfy = __fmaf_rn( qbyte<J>( fary, K ), ayf, byf ),
				tfz = __fmaf_rn( qbyte<J>( farz, K ), azf, bzf );
	const float tn = fmaxf( max3f( tnx, tny, tnz ), r.tmin );
	const float tf = fminf( min3f( tfx, tfy, tfz ), tcull );
	if (tn <= tf)
	{
		const uint32_t bits = (childBits4 >> (8 * J)) & 0xffu, idx = (bitIndex4 >> (8 * J)) & 0xffu;
		hitmask |= bits << idx;
	}
}

__device__ __forceinline__ uint32_t intersectNode( const RayState& r, const float tScale, const uint32_t K, const uint4 n0, const uint4 n1, const uint4 n2,
	const uint4 n3, const uint4 n4 )
{
	const uint32_t ew = n0.w;
	const float sx = __uint_as_float( (ew & 0xffu) << 23 ), sy = __uint_as_float( ((ew >> 8) & 0xffu) << 23 ),
				sz = __uint_as_float( ((ew >> 16) & 0xffu) << 23 );
	// plane t = (p + b*s - O) * idir = F*A + B with F = 2^23 + 256 b (see qbyte): A = s*idir/256,
	// B = ((p - O) - 32768 s) * idir.  Two things keep the test conservative.  Every plane carries a guard band of at least 1/64 cell
	// (QuantiseChild; the node frame has headroom so that no plane loses it to clamping), which covers the 1/512-cell rounding of the
	// one-FFMA decode.  And the far planes are scaled by 1+2^-20: the other error, the rounding of (p - O), is a few 2^-24 of the plane
	// distance whatever the cell size, more than the guard band of a small box far from the ray origin; tn <= tf only compares the two,
	// so a one-sided margin of twice the size serves both.  (A far plane behind the origin moves the wrong way: the box is culled anyway.)
	const float WIDEN = 1.0f + 9.53674316e-7f;
#if B200_HALF_DECODE
	// F = 1024 + b: A = s*idir, B = ((p - O) - 1024 s) * idir
	const float ax = sx * r.ix, ay = sy * r.iy, az = sz * r.iz;
	const float bx = __fmaf_rn( -1024.f, sx, __uint_as_float( n0.x ) - r.ox ) * r.ix;
	const float by = __fmaf_rn( -1024.f, sy, __uint_as_float( n0.y ) - r.oy ) * r.iy;
	const float bz = __fmaf_rn( -1024.f, sz, __uint_as_float( n0.z ) - r.oz ) * r.iz;
#else
	const float ax = (sx * 0.00390625f) * r.ix, ay = (sy * 0.00390625f) * r.iy, az = (sz * 0.00390625f) * r.iz;
	const float bx = __fmaf_rn( -32768.f, sx, __uint_as_float( n0.x ) - r.ox ) * r.ix;
	const float by = __fmaf_rn( -32768.f, sy, __uint_as_float( n0.y ) - r.oy ) * r.iy;
	const float bz = __fmaf_rn( -32768.f, sz, __uint_as_float( n0.z ) - r.oz ) * r.iz;
#endif
	const float axf = ax * WIDEN, ayf = ay * WIDEN, azf = az * WIDEN, bxf = bx * WIDEN, byf = by * WIDEN, bzf = bz * WIDEN;
	const bool nx = r.dx < 0, ny = r.dy < 0, nz = r.dz < 0;
	// Culling against the current hit distance leaves a margin of 2^-18 t.  A child's entry distance is computed from (p - O) * idir in
	// floats: its error (a few ulp of t) can exceed the quantisation guard band of a small, distant box, and a box entered at exactly the
	// hit distance (a triangle lying in its near face) could then be culled or not depending on whether a hit one ulp farther had
	// already been found -- i.e. on the order in which the warp happened to test triangles.  With the margin the set of triangles
	// that can win never depends on that order (the wave's output is reproducible; see also the tie rule in intersectTriangle).
	// (tScale: 1 + 2^-18 inside an instance; at the top level additionally the factor between world-space distances and the object-space
	// units r.t is kept in, TraceArgs::tlasTScale)
	const float tcull = r.t * tScale;
	uint32_t hitmask = 0;
#pragma unroll
	for (int i = 0; i < 2; i++)
	{
		const uint32_t meta4 = i ? n1.w : n1.z;
		const uint32_t qlox = i ? n2.y : n2.x, qloy = i ? n2.w : n2.z, qloz = i ? n3.y : n3.x;
		const uint32_t qhix = i ? n3.w : n3.z, qhiy = i ? n4.y : n4.x, qhiz = i ? n4.w : n4.z;
		const uint32_t nearx = nx ? qhix : qlox, farx = nx ? qlox : qhix;
		const uint32_t neary = ny ? qhiy : qloy, fary = ny ? qloy : qhiy;
		const uint32_t nearz = nz ? qhiz : qloz, farz = nz ? qloz : qhiz;
		const uint32_t isInner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
		const uint32_t innerMask4 = sign_extend_s8x4( isInner4 << 3 );
		const uint32_t bitIndex4 = (meta4 ^ (r.octinv4 & innerMask4)) & 0x1f1f1f1fu;
		const uint32_t childBits4 = (meta4 >> 5) & 0x07070707u;
#if B200_HALF_DECODE
		float fnx[4], fny[4], fnz[4], ffx[4], ffy[4], ffz[4];
		qpair<0>( nearx, K, fnx[0], fnx[1] ), qpair<1>( nearx, K, fnx[2], fnx[3] ), qpair<0>( neary, K, fny[0], fny[1] ), qpair<1>( neary, K, fny[2], fny[3] );
		qpair<0>( nearz, K, fnz[0], fnz[1] ), qpair<1>( nearz, K, fnz[2], fnz[3] ), qpair<0>( farx, K, ffx[0], ffx[1] ), qpair<1>( farx, K, ffx[2], ffx[3] );
		qpair<0>( fary, K, ffy[0], ffy[1] ), qpair<1>( fary, K, ffy[2], ffy[3] ), qpair<0>( farz, K, ffz[0], ffz[1] ), qpair<1>( farz, K, ffz[2], ffz[3] );
#pragma unroll
		for (int j = 0; j < 4; j++)
		{
			const float tn = fmaxf( max3f( __fmaf_rn( fnx[j], ax, bx ), __fmaf_rn( fny[j], ay, by ), __fmaf_rn( fnz[j], az, bz ) ), r.tmin );
			const float tf = fminf( min3f( __fmaf_rn( ffx[j], axf, bxf ), __fmaf_rn( ffy[j], ayf, byf ), __fmaf_rn( ffz[j], azf, bzf ) ), tcull );
			if (tn <= tf)
			{
				const uint32_t bits = (childBits4 >> (8 * j)) & 0xffu, idx = (bitIndex4 >> (8 * j)) & 0xffu;
				hitmask |= bits << idx;
			}
		}
#else
		testChild<0>( r, tcull, K, nearx, neary, nearz, farx, fary, farz, ax, ay, az, bx, by, bz, axf, ayf, azf, bxf, byf, bzf, childBits4, bitIndex4, hitmask );
		testChild<1>( r, tcull, K, nearx, neary, nearz, farx, fary, farz, ax, ay, az, bx, by, bz, axf, ayf, azf, bxf, byf, bzf, childBits4, bitIndex4, hitmask );
		testChild<2>( r, tcull, K, nearx, neary, nearz, farx, fary, farz, ax, ay, az, bx, by, bz, axf, ayf, azf, bxf, byf, bzf, childBits4, bitIndex4, hitmask );
		testChild<3>( r, tcull, K, nearx, neary, nearz, farx, fary, farz, ax, ay, az, bx, by, bz, axf, ayf, azf, bxf, byf, bzf, childBits4, bitIndex4, hitmask );
#endif
	}
	return hitmask;
}

// ---- the wave kernel ------------------------------------------------------------------------------------
// Lanes own rays and walk the wide BVH; leaf triangles are NOT tested by the lane that found them.  Each lane
// appends (owner lane, triangle) pairs to a per-warp ring in shared memory and the warp tests 32 of them at a
// time, one per lane, against the owners' rays mirrored in shared memory (a lane-divergent `while (triangles)`
// loop ran with ~3 of 32 lanes active on incoherent rays).  Results go back through two ATOMS.MIN per hit:
// smallest t, then (exact ties) the smallest leaf-order triangle index -- see intersectTriangle.
constexpr int WARPS_PER_BLOCK = BLOCK_THREADS / 32;
#ifndef B200_REFILL_BELOW
#define B200_REFILL_BELOW 22
#endif
#ifndef B200_PARTIAL_BELOW
#define B200_PARTIAL_BELOW 12
#endif
#ifndef B200_MIN_BLOCKS
#define B200_MIN_BLOCKS 6
#endif
#ifndef B200_STREAM_RAYS
#define B200_STREAM_RAYS 1 // rays are fetched and results stored with the evict-first policy (ld.global.cs / st.global.cs): they are touched once, and
                            // on scenes larger than L2 (C3: 581 MB) every line they do not claim is a node or triangle line that stays
#endif
#ifndef B200_FLUSH_AT
#define B200_FLUSH_AT 24 // test the pending triangles as soon as 24 wait (32 per round at most): hit distances shrink earlier, +2 % over 32
#endif
#ifndef B200_STEAL
#define B200_STEAL 0 // 1: at the end of a wave, lanes without a ray take over stack entries of the rays still walking (one-instance kernels; see "work
                     // donation" in the pass).  Correct (the parity suite passes with it) and measured: the isolated 1080p bounce waves gain 5 %
                     // (2185 -> 2299 Mrays/s), frames gain nothing (their chains already fill each other's tails) and the large waves lose 2.7 % to
                     // the extra state in the pass (4K frame 127.0 -> 130.4 ms).  Off.
#endif
#ifndef B200_DRAIN_PREFETCH
#define B200_DRAIN_PREFETCH 0 // n > 0: a warp that has run out of rays and has fewer than n lanes active prefetches its next node (see B' in the pass)
#endif
#ifndef B200_EAGER_BELOW
#define B200_EAGER_BELOW 8 // test pending triangles every pass when fewer lanes than this have node work
#endif
constexpr int RING_CAP = 64;                          // leaf records per warp ring (a pass adds at most 32)
constexpr int REFILL_BELOW = B200_REFILL_BELOW;      // leave the traversal loop to fetch rays when fewer lanes are active
constexpr int PARTIAL_BELOW = B200_PARTIAL_BELOW;    // flush a partially filled ring for waiting lanes when fewer lanes have node work

// SINGLE: the scene is ONE instance (record 0).  Rays are taken into its space when they are fetched and never come back to
// the top level: no top-level node visit, no instance step in the pass, no world-space copy of the ray, one-level stack logic.
template <int ISECT, bool ANYHIT, bool COUNT, bool SINGLE>
__global__ void __launch_bounds__( BLOCK_THREADS, B200_MIN_BLOCKS ) traceKernel( const TraceArgs a )
{
	__shared__ uint2 sstack[STACK_SMEM * BLOCK_THREADS];
	__shared__ float4 sRayO[BLOCK_THREADS];          // current-space ray of each lane: (ox,oy,oz,tmin)
	__shared__ float4 sRayD[BLOCK_THREADS];          //                                  (dx,dy,dz,t); .w is ATOMS.MIN'ed as uint bits
	__shared__ uint32_t sSlot[BLOCK_THREADS];        // winner of a test round: (leaf-order triangle index << 5 | tester lane), 0xffffffff = none
	__shared__ uint32_t sBestTri[BLOCK_THREADS];     // leaf-order index of the triangle of each lane's current hit (tie rule, see intersectTriangle)
	__shared__ uint32_t sHelp[B200_STEAL && SINGLE ? BLOCK_THREADS : 1]; // work donation: helpers still walking for this lane's ray; bit 31: the ray is over (any-hit), helpers stop
	__shared__ float4 sWorld[SINGLE ? 1 : BLOCK_THREADS]; // world-space (1/d, octant) of each lane's ray, restored when it leaves an instance
	__shared__ float4 sWorldO[SINGLE ? 1 : BLOCK_THREADS], sWorldD[SINGLE ? 1 : BLOCK_THREADS]; // ... and the ray itself (re-reading the ray planes there cost 2 %)
	__shared__ uint2 sRing[WARPS_PER_BLOCK][RING_CAP]; // leaf records: (first triangle | owner lane << 27, 24-bit triangle mask)
	uint2 lstack[STACK_LOCAL];
	const uint32_t lane = threadIdx.x & 31, warpBase = threadIdx.x & ~31u;
	uint2* const ring = sRing[threadIdx.x >> 5];
	const uint32_t total = a.countPtr ? __ldg( a.countPtr ) : a.count;
	const uint4* __restrict__ nodes = a.nodes;
	const float4* __restrict__ tris = a.tris;
	const uint32_t K = a.magic; // see qbyte / qpair; set by the launchers
	const float tScaleTop = a.tlasTScale * (1.0f + 3.81469727e-6f);
	const float4 singleInfo = SINGLE ? __ldg( a.instances + 4 ) : make_float4( 0, 0, 0, 0 ); // (BLAS root, instance id, ..) of record 0

	RayState r = {};
	uint32_t rayIdx = 0, hitPrim = 0, hitInst = 0, hitBary = 0, curInst = 0, nNodes = 0, nTris = 0;
	uint2 ngroup = make_uint2( 0, 0 ), tgroup = make_uint2( 0, 0 );
	int sp = 0, instBase = -1;
	uint32_t qhead = 0, qtail = 0;     // ring sequence numbers of the leaf records (warp-uniform)
	uint32_t pendingTris = 0;          // triangles in the ring (warp-uniform)
	uint32_t lastQueued = 0xffffffffu; // sequence number of this lane's newest record
	bool active = false, exhausted = false, terminated = false, finished = false;
	constexpr bool STEAL = B200_STEAL && SINGLE && !COUNT;
	bool helping = false;            // this lane walks a donated part of another lane's ray
	uint32_t ownerLane = lane;       // the lane whose ray this lane's triangle records belong to (its own, unless helping)
	sSlot[threadIdx.x] = 0xffffffffu;
	if (STEAL) sHelp[threadIdx.x] = 0;
	if (a.fetchCounterNext && blockIdx.x == 0 && threadIdx.x == 0) *a.fetchCounterNext = 0;

	auto push = [&]( const uint2 g ) {
		if (sp < STACK_SMEM) sstack[sp * BLOCK_THREADS + threadIdx.x] = g; else lstack[sp - STACK_SMEM] = g;
		sp++;
	};
	auto pop = [&]() -> uint2 {
		sp--;
		return sp < STACK_SMEM ? sstack[sp * BLOCK_THREADS + threadIdx.x] : lstack[sp - STACK_SMEM];
	};
	auto mirrorRay = [&]() {
		sRayO[threadIdx.x] = make_float4( r.ox, r.oy, r.oz, r.tmin );
		sRayD[threadIdx.x] = make_float4( r.dx, r.dy, r.dz, r.t );
	};
	// One round: the first min(32, pending) triangles of the ring, one per lane; every lane of the warp calls this (converged).
	// Lane r loads record r, a warp scan of the triangle counts gives each record its range, tester lane L finds its record
	// with a 5-step search over the scanned counts and takes that record's (L - start)-th triangle.
	auto flush = [&]() {
		const uint32_t nrec = min( qtail - qhead, 32u );
		const uint2 rec = lane < nrec ? ring[(qhead + lane) & (RING_CAP - 1)] : make_uint2( 0u, 0u );
		const uint32_t cnt = __popc( rec.y );
		uint32_t incl = cnt;
#pragma unroll
		for (int d = 1; d < 32; d <<= 1)
		{
			const uint32_t up = __shfl_up_sync( 0xffffffffu, incl, d );
			if (lane >= d) incl += up;
		}
		const uint32_t start = incl - cnt;
		// the 32 records at hand hold all pending triangles, or (more than 32 records, one triangle or more each) at least 32
		const uint32_t take = min( pendingTris, 32u );
		// every record holds >= 1 triangle, so record starts are distinct: bit s of `starts` = a record starts at triangle s (bit 0,
		// always the first record's, carries "a record ends exactly at `take`" through the reduction instead);
		// the record of tester lane L is the number of starts at or below L, minus one
		const uint32_t red = __reduce_or_sync( 0xffffffffu, lane < nrec ? ((start - 1u < 31u ? 1u << start : 0u) | (incl == take ? 1u : 0u)) : 0u );
		const uint32_t starts = red | 1u;
		const uint32_t ri = min( (uint32_t)__popc( starts & (0xffffffffu >> (31 - lane)) ) - 1u, 31u );
		const uint32_t rx = __shfl_sync( 0xffffffffu, rec.x, ri );
		const uint32_t pk = __shfl_sync( 0xffffffffu, rec.y | (start << 24), ri ); // 24-bit mask | start (< 32 for every record a tester can have)
		const uint32_t ry = pk & 0x00ffffffu, rstart = pk >> 24;
		const bool mine = lane < take;
		bool hit = false;
		float tt = 0, u = 0, v = 0;
		uint32_t owner = 0, prim = 0, tri = 0;
		if (mine)
		{
			// my triangle is the k-th highest set bit of the record's mask, k = lane - rstart.  Walk to it from the nearer end, in
			// a form where one step is `m &= m - 1` (two dependent instructions): from the top on the bit-reversed mask, else from
			// the bottom.  Records of up to three triangles need at most one step; the loop runs for the warp's largest count.
			const uint32_t k = lane - rstart, cnt = __popc( ry );
			const bool fromTop = 2 * k < cnt;
			uint32_t m = fromTop ? __brev( ry ) : ry;
			for (uint32_t j = fromTop ? k : cnt - 1 - k; j; j--) m &= m - 1;
			const uint32_t low = __ffs( m ) - 1;
			const uint32_t bit = fromTop ? 31 - low : low;
			owner = warpBase + (rx >> 27);
			tri = (rx & 0x07ffffffu) + bit;
			const float4* tp = tris + (size_t)tri * 3;
			const float4 v0 = __ldg( tp ), e1 = __ldg( tp + 1 ), e2 = __ldg( tp + 2 );
			const float4 O = sRayO[owner], D = sRayD[owner];
			prim = __float_as_uint( v0.w );
			hit = intersectTriangle<ISECT>( O, D, v0, e1, e2, tri, &sBestTri[owner], tt, u, v );
		}
		// retire consumed records; the record that straddles the 32-triangle boundary keeps its untested triangles
		if (lane < nrec && incl > take && incl - cnt < take)
		{
			ring[(qhead + lane) & (RING_CAP - 1)].y = dropHighest( rec.y, take - (incl - cnt) );
		}
		// records used up = records that end at or before `take`: one per start at 1 .. take-1, plus the one that ends exactly there
		qhead += __popc( starts >> 1 ) + (red & 1u);
		pendingTris -= take;
		const uint32_t hitMask = __ballot_sync( 0xffffffffu, hit ); // also orders the mirror reads before the atomics
		if (hitMask)
		{
			uint32_t* const tbest = reinterpret_cast<uint32_t*>( &sRayD[owner].w );
			if (hit) atomicMin( tbest, __float_as_uint( tt ) );
			__syncwarp();
			if (hit && *tbest == __float_as_uint( tt )) atomicMin( &sSlot[owner], (tri << 5) | lane ); // (triangle indices stay below 2^27)
			__syncwarp();
			const uint32_t slot = sSlot[threadIdx.x];
			const bool changed = slot != 0xffffffffu;
			const uint32_t src = changed ? (slot & 31u) : lane;
			const uint32_t bary = (uint32_t)(65535.0f * u) + ((uint32_t)(65535.0f * v) << 16);
			const uint32_t wPrim = __shfl_sync( 0xffffffffu, prim, src ), wBary = __shfl_sync( 0xffffffffu, bary, src );
			const float wT = __shfl_sync( 0xffffffffu, tt, src );
			if (changed)
			{
				sSlot[threadIdx.x] = 0xffffffffu, sBestTri[threadIdx.x] = slot >> 5;
				r.t = wT, hitPrim = wPrim, hitBary = wBary;
				if (!SINGLE) hitInst = curInst;
				if (ANYHIT) terminated = true, ngroup.y = 0, tgroup.y = 0;
				if (ANYHIT && STEAL) atomicOr( &sHelp[threadIdx.x], 0x80000000u );
			}
		}
		__syncwarp();
		if (STEAL && helping)
		{
			// a helper culls against the distance of the ray it works for (kept in the mirror by the atomics above)
			r.t = sRayD[warpBase + ownerLane].w;
			if (ANYHIT && (sHelp[warpBase + ownerLane] & 0x80000000u)) ngroup.y = 0, tgroup.y = 0, sp = 0;
		}
	};

	auto writeResult = [&]() {
			if (ANYHIT)
			{
				if (a.accumulator)
				{
					// fused connect: an unoccluded shadow ray delivers its potential contribution
					if (hitPrim == 0xffffffffu)
					{
						const float4 E = __ldg( a.potentials + rayIdx );
						float* px = reinterpret_cast<float*>( a.accumulator + __float_as_uint( E.w ) );
						atomicAdd( px, E.x ), atomicAdd( px + 1, E.y ), atomicAdd( px + 2, E.z );
					}
				}
				else if (hitPrim != 0xffffffffu) atomicOr( a.occluded + (rayIdx >> 5), 1u << (rayIdx & 31) );
			}
			else
			{
				const bool hit = hitPrim != 0xffffffffu;
				const float4 record = make_float4( __uint_as_float( hit ? hitBary : 0u ), __uint_as_float( hit ? (SINGLE ? __float_as_uint( singleInfo.y ) : hitInst) : 0u ),
					__uint_as_float( hitPrim ), hit ? r.t : 1e34f );
#if B200_STREAM_RAYS
				__stcs( a.hits + rayIdx, record );
#else
				a.hits[rayIdx] = record;
#endif
			}
			if (COUNT) a.stepNodes[rayIdx] = nNodes, a.stepTris[rayIdx] = nTris;
			finished = false;
	};

	while (true)
	{
		// ---- results of the lanes that retired since the last fetch ------------------------------------
		if (finished) writeResult();
		// ---- ray fetch: idle lanes claim consecutive rays with one aggregated atomic ------------------
		const uint32_t idleMask = __ballot_sync( 0xffffffffu, !active );
		if (idleMask && !exhausted)
		{
			const uint32_t n = __popc( idleMask ), leader = __ffs( idleMask ) - 1;
			uint32_t base = 0;
			if (lane == leader) base = atomicAdd( a.fetchCounter, n );
			base = __shfl_sync( 0xffffffffu, base, leader );
			if (!active)
			{
				rayIdx = base + __popc( idleMask & lanemask_lt() );
				if (rayIdx < total)
				{
#if B200_STREAM_RAYS
					const float4 O4 = __ldcs( (a.hostO ? a.hostO : a.rayO) + rayIdx ), D4 = __ldcs( (a.hostO ? a.hostD : a.rayD) + rayIdx );
#else
					const float4 O4 = __ldg( (a.hostO ? a.hostO : a.rayO) + rayIdx ), D4 = __ldg( (a.hostO ? a.hostD : a.rayD) + rayIdx );
#endif
					if (a.hostO) a.stageO[rayIdx] = O4, a.stageD[rayIdx] = D4; // host-buffer wave: the upload happens here
					r.ox = O4.x, r.oy = O4.y, r.oz = O4.z, r.tmin = O4.w;
					r.dx = D4.x, r.dy = D4.y, r.dz = D4.z, r.t = D4.w;
					if (SINGLE)
					{
						// straight into the instance (reference arithmetic of the instance entry).  The top-level node, whose only box is
						// this instance's, is not visited: the bottom-level root test rejects what it would have rejected
						enterInstance<ISECT>( r, a.instances );
						mirrorRay();
						ngroup = make_uint2( __float_as_uint( singleInfo.x ), 0x80000000u ), tgroup = make_uint2( 0, 0 );
					}
					else
					{
						setupSlab( r );
						sWorld[threadIdx.x] = make_float4( r.ix, r.iy, r.iz, __uint_as_float( r.octinv4 ) );
						sWorldO[threadIdx.x] = O4, sWorldD[threadIdx.x] = D4;
						ngroup = make_uint2( a.tlasRoot, 0x80000000u ), tgroup = make_uint2( 0, 0 );
					}
					sp = 0, instBase = -1, curInst = NO_INSTANCE;
					hitPrim = 0xffffffffu, hitInst = 0, hitBary = 0, nNodes = nTris = 0;
					lastQueued = qhead - 1, terminated = false;
					sBestTri[threadIdx.x] = 0xffffffffu;
					if (STEAL) sHelp[threadIdx.x] = 0;
					active = true;
				}
			}
			if (base + n >= total) exhausted = true;
		}
		if (!__any_sync( 0xffffffffu, active )) break;

		// ---- traversal: every pass = one node step per lane, then queue / flush triangle pairs ------------
		uint32_t votes = 0x00202000u; // "32 lanes active, 32 with work, none waiting": the first pass of a round decides nothing
		while (true)
		{
			// Work donation.  When the wave has no rays left to hand out, its end is a few long rays walking alone, each pass one dependent
			// chain at memory latency, while the other lanes of their warps idle (T(n) = 38 us + n / 3.27 Grays/s: the constant).  An idle lane
			// then takes the newest stack entry of a lane that still walks and traverses that subtree FOR the same ray: the ray is in the
			// shared-memory mirror anyway (the triangle testers read it there), hits come home through the mirror's atomicMin and the owner's
			// slot word exactly as a tester's do, and the owner retires only when its helpers have.  Which lane finds a hit never shows in
			// the result (tie rule, cull margin).  One-instance kernels only: a helper works in the instance's space, which the owner of a
			// two-level ray leaves and re-enters.
			if (STEAL && exhausted)
			{
				if (ANYHIT && helping && (sHelp[warpBase + ownerLane] & 0x80000000u)) ngroup.y = 0, tgroup.y = 0, sp = 0; // the shadow ray has found its occluder
				if (finished) writeResult(); // (a retired lane's result still sits in the registers a helper needs)
				const uint32_t idleMask = __ballot_sync( 0xffffffffu, !active );
				if (idleMask)
				{
					const bool canGive = active && !terminated && sp > 0;
					const uint32_t donorMask = __ballot_sync( 0xffffffffu, canGive );
					if (donorMask)
					{
						const uint32_t pairs = min( __popc( idleMask ), __popc( donorMask ) );
						uint2 g = make_uint2( 0, 0 );
						if (canGive && (uint32_t)__popc( donorMask & lanemask_lt() ) < pairs) g = pop();
						const uint32_t rank = __popc( idleMask & lanemask_lt() );
						const bool take = !active && rank < pairs;
						const uint32_t from = take ? __fns( donorMask, 0, rank + 1 ) : lane;
						const uint32_t gx = __shfl_sync( 0xffffffffu, g.x, from ), gy = __shfl_sync( 0xffffffffu, g.y, from );
						const uint32_t forLane = __shfl_sync( 0xffffffffu, ownerLane, from );
						if (take)
						{
							ownerLane = forLane, helping = true;
							const float4 O4 = sRayO[warpBase + ownerLane], D4 = sRayD[warpBase + ownerLane];
							r.ox = O4.x, r.oy = O4.y, r.oz = O4.z, r.tmin = O4.w, r.dx = D4.x, r.dy = D4.y, r.dz = D4.z, r.t = D4.w;
							setupSlab( r );
							ngroup = make_uint2( gx, gy ), tgroup = make_uint2( 0, 0 );
							sp = 0, lastQueued = qhead - 1, terminated = false, active = true;
							atomicAdd( &sHelp[warpBase + ownerLane], 1u );
						}
						__syncwarp();
					}
				}
			}
			// A. top level: a hit leaf is an instance -- remember what is left here, then descend into its BLAS (its root is visited
			//    by step B of the same pass)
			if (!SINGLE && active && instBase < 0 && tgroup.y)
			{
				const uint32_t bit = 31 - __clz( tgroup.y );
				tgroup.y &= ~(1u << bit);
				const uint32_t idx = tgroup.x + bit;
				if (ngroup.y > 0x00ffffffu) push( ngroup );
				if (tgroup.y) push( tgroup );
				const float4* rec = a.instances + (size_t)idx * 6;
				const float4 info = __ldg( rec + 4 );
				enterInstance<ISECT>( r, rec );
				mirrorRay();
				curInst = __float_as_uint( info.y );
				instBase = sp;
				ngroup = make_uint2( __float_as_uint( info.x ), 0x80000000u );
				tgroup = make_uint2( 0, 0 );
			}
			// B. node step (not while instance leaves of the previous step are still waiting to be entered)
			if (active && ngroup.y > 0x00ffffffu && (SINGLE || tgroup.y == 0))
			{
				const uint32_t hits = ngroup.y, imask = hits & 0xffu;
				const uint32_t bit = 31 - __clz( hits );
				ngroup.y = hits & ~(1u << bit);
				const uint32_t nodeBase = ngroup.x;
				if (ngroup.y > 0x00ffffffu) push( ngroup );
				const uint32_t slot = (bit - 24) ^ (r.octinv4 & 0xffu);
				const uint32_t rel = __popc( imask & ~(0xffffffffu << slot) );
				const uint4* np = nodes + (size_t)(nodeBase + rel) * 5;
				const uint4 n0 = __ldg( np ), n1 = __ldg( np + 1 ), n2 = __ldg( np + 2 ), n3 = __ldg( np + 3 ), n4 = __ldg( np + 4 );
				if (COUNT) nNodes++;
				const uint32_t hm = intersectNode( r, (!SINGLE && instBase < 0) ? tScaleTop : (1.0f + 3.81469727e-6f), K, n0, n1, n2, n3, n4 );
				ngroup = make_uint2( n1.x, (hm & 0xff000000u) | (n0.w >> 24) );
				tgroup = make_uint2( n1.y, hm & 0x00ffffffu );
			}
#if B200_DRAIN_PREFETCH
			// B'. the end of a wave: a few rays walk alone and every pass is one dependent chain -- node fetch, test, triangle fetch, test --
			//     at L2 latency.  The node of the NEXT pass is already known here (the nearest inner hit, else the top of the stack): ask for
			//     its lines now, so that the fetch runs beside the triangle round of this pass instead of behind it.  Only when the warp has
			//     run out of rays and of parallelism (the prefetch costs issue slots, which the bulk of the wave does not have to spare).
			if (exhausted && (votes >> 16) < B200_DRAIN_PREFETCH && active)
			{
				uint2 g = ngroup;
				if (g.y <= 0x00ffffffu && sp > (SINGLE ? 0 : (instBase < 0 ? 0 : instBase)))
					g = sp - 1 < STACK_SMEM ? sstack[(sp - 1) * BLOCK_THREADS + threadIdx.x] : lstack[sp - 1 - STACK_SMEM];
				if (g.y > 0x00ffffffu)
				{
					const uint32_t bit = 31 - __clz( g.y ), slot = (bit - 24) ^ (r.octinv4 & 0xffu);
					const uint4* next = nodes + (size_t)(g.x + __popc( g.y & 0xffu & ~(0xffffffffu << slot) )) * 5;
					asm volatile( "prefetch.global.L1 [%0];" ::"l"( next ) );
					asm volatile( "prefetch.global.L1 [%0];" ::"l"( next + 4 ) );
				}
			}
#endif
			// C. bottom level: a lane that hit leaf triangles appends ONE record (first triangle, mask, owner) to the warp's
			//    ring; as soon as 32 triangles are pending the warp tests them, one per lane
			{
				const bool has = active && (SINGLE || instBase >= 0) && tgroup.y != 0;
				const uint32_t m = __ballot_sync( 0xffffffffu, has );
				if (m)
				{
					if (has)
					{
						lastQueued = qtail + __popc( m & lanemask_lt() );
						ring[lastQueued & (RING_CAP - 1)] = make_uint2( tgroup.x | ((STEAL ? ownerLane : lane) << 27), tgroup.y );
						if (COUNT) nTris += __popc( tgroup.y );
					}
					pendingTris += __reduce_add_sync( 0xffffffffu, has ? (uint32_t)__popc( tgroup.y ) : 0u );
					if (has) tgroup.y = 0;
					qtail += __popc( m );
					__syncwarp();
					while (pendingTris >= B200_FLUSH_AT) flush();
				}
			}
			// D. lanes without node work: leave the instance / pop / retire (only once their queued pairs are tested).  A retired
			//    lane keeps its result in registers; it is written when the warp next fetches rays (see the top of the outer loop),
			//    several lanes at a time, so that the per-pass path holds no result stores.  The world-space ray is only restored
			//    when top-level work is left on the stack.
			bool waiting = false;
			if (active && ngroup.y <= 0x00ffffffu && (SINGLE || tgroup.y == 0)) // (SINGLE: step C has always taken the triangles)
			{
				const bool pending = (int32_t)(lastQueued - qhead) >= 0;
				const bool leaving = terminated || sp == (SINGLE ? 0 : instBase);
				if (leaving && pending) waiting = true;
				else if (STEAL && helping && (terminated || sp == 0))
				{
					// a helper is done when its subtree is and its records are tested: it reports back and is idle again
					atomicSub( &sHelp[warpBase + ownerLane], 1u );
					active = false, helping = false, ownerLane = lane;
				}
				else if (STEAL && (terminated || sp == 0) && (sHelp[threadIdx.x] & 0x7fffffffu)) waiting = true; // the ray's helpers are still walking
				else if (terminated || sp == 0) active = false, finished = true;
				else
				{
					if (!SINGLE && leaving)
					{
						// bottom level exhausted: back to the world-space ray (t is kept, as in bvh.cpp:612-619)
						const float4 O4 = sWorldO[threadIdx.x], D4 = sWorldD[threadIdx.x];
						const float4 W = sWorld[threadIdx.x];
						r.ox = O4.x, r.oy = O4.y, r.oz = O4.z, r.dx = D4.x, r.dy = D4.y, r.dz = D4.z;
						r.ix = W.x, r.iy = W.y, r.iz = W.z, r.octinv4 = __float_as_uint( W.w );
						instBase = -1, curInst = NO_INSTANCE;
					}
					const uint2 g = pop();
					if (SINGLE || g.y > 0x00ffffffu) ngroup = g; else tgroup = g, ngroup = make_uint2( 0, 0 );
				}
			}
			// E. one warp reduction counts the waiting lanes, the lanes with work and the active lanes (a byte each).  The decisions
			//    below use the counts of the PREVIOUS pass: the reduction of this pass is issued here and first read a whole pass
			//    later, so its latency is off the pass's dependent chain; acting one pass late costs nothing (a refill or a partial
			//    flush starts one pass later, and the loop makes one idle pass before it ends).
			const uint32_t prev = votes;
			votes = __reduce_add_sync( 0xffffffffu, (waiting ? 1u : 0u) | ((active && (ngroup.y > 0x00ffffffu || tgroup.y)) ? 0x100u : 0u) |
				(active ? 0x10000u : 0u) );
			//    flush a partially filled ring when lanes wait on it and little node work is left in the warp
			//    ... or, whoever waits, when so few lanes have node work that the warp is near the end of its rays: their hit distances
			//    then shrink pass by pass instead of once per 32 triangles, and the last, longest rays cull more
			const uint32_t nWork = (prev >> 8) & 0xffu;
			if (pendingTris && (nWork < B200_EAGER_BELOW || ((prev & 0xffu) && nWork < PARTIAL_BELOW))) while (pendingTris) flush();
			// F. too few lanes left in this warp and rays remain: go back and refill the idle lanes
			const uint32_t nActive = prev >> 16;
			if (!nActive || (!exhausted && nActive < REFILL_BELOW)) break;
		}
	}
}

} // namespace

#ifndef B200_SMEM_CARVEOUT
#define B200_SMEM_CARVEOUT -1 // percent of the 228 KB kept as shared memory (-1: the driver's choice); the rest of the 256 KB is L1
#endif
namespace {
typedef void ( *TraceKernel )( const TraceArgs );
// resident CTAs per SM of a kernel instance (asked once per instance)
int blocksPerSM( TraceKernel kernel )
{
	static TraceKernel known[16];
	static int blocks[16], count = 0;
	for (int i = 0; i < count; i++) if (known[i] == kernel) return blocks[i];
	int n = 0;
	if (B200_SMEM_CARVEOUT >= 0) cudaFuncSetAttribute( kernel, cudaFuncAttributePreferredSharedMemoryCarveout, B200_SMEM_CARVEOUT );
	cudaOccupancyMaxActiveBlocksPerMultiprocessor( &n, kernel, BLOCK_THREADS, 0 );
	if (n < 1) n = 1;
	if (count < 16) known[count] = kernel, blocks[count] = n, count++;
	return n;
}
template <bool ANYHIT> TraceKernel pick( IsectMode mode, bool countSteps, bool single )
{
	if (mode == ISECT_REFERENCE)
	{
		if (countSteps) return single ? traceKernel<ISECT_REFERENCE, ANYHIT, true, true> : traceKernel<ISECT_REFERENCE, ANYHIT, true, false>;
		return single ? traceKernel<ISECT_REFERENCE, ANYHIT, false, true> : traceKernel<ISECT_REFERENCE, ANYHIT, false, false>;
	}
	if (countSteps) return single ? traceKernel<ISECT_WATERTIGHT, ANYHIT, true, true> : traceKernel<ISECT_WATERTIGHT, ANYHIT, true, false>;
	return single ? traceKernel<ISECT_WATERTIGHT, ANYHIT, false, true> : traceKernel<ISECT_WATERTIGHT, ANYHIT, false, false>;
}
} // namespace

cudaError_t LaunchExtend( const TraceArgs& args, IsectMode mode, bool countSteps, int smCount, cudaStream_t s )
{
	TraceArgs a = args;
	a.magic = B200_HALF_DECODE ? 0x64646464u : 0x4B000000u;
	cudaError_t e = a.fetchCounterNext ? cudaSuccess : cudaMemsetAsync( a.fetchCounter, 0, 4, s );
	if (e != cudaSuccess) return e;
	const TraceKernel k = pick<false>( mode, countSteps, a.singleInstance != 0 );
	k<<<smCount * blocksPerSM( k ), BLOCK_THREADS, 0, s>>>( a );
	return cudaGetLastError();
}

// ---- the extend wave without any allocation (b200_trace_closest_noalloc) ---------------------------------------------------------
// No counterpart in the reference; added for callers that cannot call cudaMalloc (code under
// CUDA graph capture, tight serving loops, harnesses that forbid it).  The wave kernel's scratch -- the persistent-thread fetch counter
// and the device staging of the ray planes and the hit records of a host-buffer wave -- lives in ONE file-scope device array of a
// fixed size instead of the core's growable cudaMalloc'ed buffers (rendercore.cpp::hostWave).  Every GPU that loads the library holds
// the array until exit; every call shares it and it keeps the last call's contents, so the fetch counter is cleared by every call and
// two calls that overlap on the GPU (different streams, different cores on one device) race.
// Layout: [0, 256) fetch counter word | O4 plane | D4 plane | hit records, each NOALLOC_MAX_RAYS x 16 bytes.
static __device__ __align__( 256 ) unsigned char g_noAllocScratch[256 + (size_t)NOALLOC_MAX_RAYS * 48];

// In host code the array's NAME is a host-side stand-in, not a device address: the address comes from cudaGetSymbolAddress, asked once
// per device with that device current.
static cudaError_t noAllocScratchAddress( int device, unsigned char** out )
{
	constexpr int MAX_DEVICES = 64;
	static unsigned char* address[MAX_DEVICES] = {};
	if (device < 0 || device >= MAX_DEVICES) return cudaErrorInvalidDevice;
	if (!address[device])
	{
		void* p = nullptr;
		const cudaError_t e = cudaGetSymbolAddress( &p, g_noAllocScratch );
		if (e != cudaSuccess) return e;
		address[device] = (unsigned char*)p;
	}
	*out = address[device];
	return cudaSuccess;
}

cudaError_t TraceClosestNoAlloc( const TraceArgs& base, IsectMode mode, int smCount, int device, const float* hostO4, const float* hostD4, uint32_t n, float* hostHits,
	cudaStream_t s )
{
	if (n > NOALLOC_MAX_RAYS) return cudaErrorInvalidValue;
	if (n == 0) return cudaSuccess;
	int previous = -1;
	cudaError_t e = cudaGetDevice( &previous );
	if (e != cudaSuccess) return e;
	if (previous != device && (e = cudaSetDevice( device )) != cudaSuccess) return e;
	unsigned char* scratch = nullptr;
	e = noAllocScratchAddress( device, &scratch );
	if (e == cudaSuccess)
	{
		float4* dO = reinterpret_cast<float4*>( scratch + 256 );
		float4* dD = dO + NOALLOC_MAX_RAYS;
		float4* dHits = dD + NOALLOC_MAX_RAYS;
		TraceArgs a = base;
		a.rayO = dO, a.rayD = dD, a.hits = dHits, a.count = n, a.countPtr = nullptr;
		a.hostO = a.hostD = nullptr, a.stageO = a.stageD = nullptr, a.stepNodes = a.stepTris = nullptr;
		a.fetchCounter = reinterpret_cast<uint32_t*>( scratch ), a.fetchCounterNext = nullptr; // LaunchExtend zeroes the counter: the array keeps the last call's value
		e = cudaMemcpyAsync( dO, hostO4, (size_t)n * 16, cudaMemcpyHostToDevice, s );
		if (e == cudaSuccess) e = cudaMemcpyAsync( dD, hostD4, (size_t)n * 16, cudaMemcpyHostToDevice, s );
		if (e == cudaSuccess) e = LaunchExtend( a, mode, false, smCount, s );
		if (e == cudaSuccess) e = cudaMemcpyAsync( hostHits, dHits, (size_t)n * 16, cudaMemcpyDeviceToHost, s );
		if (e == cudaSuccess) e = cudaStreamSynchronize( s );
	}
	if (previous != device) cudaSetDevice( previous );
	return e;
}

cudaError_t LaunchShadow( const TraceArgs& args, IsectMode mode, int smCount, cudaStream_t s )
{
	TraceArgs a = args;
	a.magic = B200_HALF_DECODE ? 0x64646464u : 0x4B000000u;
	cudaError_t e = a.fetchCounterNext ? cudaSuccess : cudaMemsetAsync( a.fetchCounter, 0, 4, s );
	if (e != cudaSuccess) return e;
	const TraceKernel k = pick<true>( mode, false, a.singleInstance != 0 );
	k<<<smCount * blocksPerSM( k ), BLOCK_THREADS, 0, s>>>( a );
	return cudaGetLastError();
}

} // namespace b200
