// filter.cu -- the SVGF-style reconstruction filter and TAA behind Setting( "filter", 1 ) / Setting( "TAA", 1 ).
//
// Replaces, for the headless B200 core, the post-process of the reference's filtering core (lib/RenderCore_Optix7Filter/rendercore.cpp:
// 827-861), whose kernels live in lib/CUDA/shared_kernel_code/finalize_shared.h:102-483:
//   prepare    (:102-176)  demodulate the direct / indirect accumulator planes by the albedo, clamp, pack both into one float4 per pixel;
//                          reproject the pixel into the previous frame (diffuse: through the previous camera; specular: diamond search
//                          over the previous world positions); blend the luminance moments with their history
//   reconstruct(:205-369)  three a-trous passes (steps 1, 2, 4) with normal / depth / albedo / luminance edge stops, separately for
//                          direct and indirect light; the first also blends with the reprojected, neighbourhood-clamped history;
//                          the last remodulates with the albedo and applies the display gamma
//   TAA        (:383-424)  Mitchell-Netravali history fetch, YCoCg variance clamp, 10 % new; then an unsharp mask (:438-462) -- or,
//                          without TAA, just the gamma removal (:472-480)
// The arithmetic follows the reference step by step -- thresholds, weights, the packing formats (5:11 fixed point shading, 10-bit
// normals, 10:11:11 albedo) and its quirks are part of what a frame looks like: the 3x3 clamp window that skips column / row 1 but
// always divides by 9, the Mitchell fetch that skips row 0, the double-precision 0.0001 in the material edge stop.  The checker is the
// reference's own kernels compiled in place (tests/test_filter_gpu.py).
// The reference writes the finished frame into a GL texture through a surface; here it goes to the core's RGBA32F frame buffer.
// All kernels are one thread per pixel over 32x8 tiles of the image: streaming stencils, bound by HBM (DESIGN.md has the bytes).
// Compiled with FMA contraction on, as the reference's own nvcc build is (the Makefile's -fmad=false is for the bit-exact kernels).
#include "kernels.h"

namespace b200 {

namespace {

struct F3 { float x, y, z; };
__device__ __forceinline__ F3 f3( const float x, const float y, const float z ) { F3 r; r.x = x, r.y = y, r.z = z; return r; }
__device__ __forceinline__ F3 f3( const float4 v ) { return f3( v.x, v.y, v.z ); }
__device__ __forceinline__ F3 operator+( const F3 a, const F3 b ) { return f3( a.x + b.x, a.y + b.y, a.z + b.z ); }
__device__ __forceinline__ F3 operator-( const F3 a, const F3 b ) { return f3( a.x - b.x, a.y - b.y, a.z - b.z ); }
__device__ __forceinline__ F3 operator*( const F3 a, const F3 b ) { return f3( a.x * b.x, a.y * b.y, a.z * b.z ); }
__device__ __forceinline__ F3 operator*( const F3 a, const float s ) { return f3( a.x * s, a.y * s, a.z * s ); }
__device__ __forceinline__ F3 operator*( const float s, const F3 a ) { return f3( s * a.x, s * a.y, s * a.z ); }
__device__ __forceinline__ float dot( const F3 a, const F3 b ) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ F3 minEach( const F3 a, const float b ) { return f3( fminf( a.x, b ), fminf( a.y, b ), fminf( a.z, b ) ); }
__device__ __forceinline__ F3 clampEach( const F3 v, const F3 lo, const F3 hi )
{
	return f3( fmaxf( lo.x, fminf( v.x, hi.x ) ), fmaxf( lo.y, fminf( v.y, hi.y ) ), fmaxf( lo.z, fminf( v.z, hi.z ) ) );
}

// ---- packing formats (lib/CUDA/shared_kernel_code/tools_shared.h) ---------------------------------------------------------------
// normal, 10 bits per axis above two flag bits (:122-136)
__device__ __forceinline__ F3 unpackNormal( const uint32_t p )
{
	return f3( (float)((p >> 2) & 1023u) * (1.0f / 511.0f) - 1, (float)((p >> 12) & 1023u) * (1.0f / 511.0f) - 1, (float)(p >> 22) * (1.0f / 511.0f) - 1 );
}
// albedo, 10:11:11 (:169-183); the `min1` flavour keeps every channel above zero for the demodulation
__device__ __forceinline__ F3 unpackAlbedo( const uint32_t c )
{
	return f3( (float)(c >> 22) * (1.0f / 1023.0f), (float)((c >> 11) & 2047u) * (1.0f / 2047.0f), (float)(c & 2047u) * (1.0f / 2047.0f) );
}
__device__ __forceinline__ F3 unpackAlbedoMin1( const uint32_t c )
{
	return f3( (float)max( 1u, c >> 22 ) * (1.0f / 1023.0f), (float)max( 1u, (c >> 11) & 2047u ) * (1.0f / 2047.0f), (float)max( 1u, c & 2047u ) * (1.0f / 2047.0f) );
}
// direct + indirect light of a pixel in one float4: 5:11 unsigned fixed point, two channels per word (:312-334)
__device__ __forceinline__ float4 packLight( const F3 a, const F3 b )
{
	const uint32_t ar = (uint32_t)(fminf( a.x, 31.999f ) * 2048.0f), ag = (uint32_t)(fminf( a.y, 31.999f ) * 2048.0f), ab = (uint32_t)(fminf( a.z, 31.999f ) * 2048.0f);
	const uint32_t br = (uint32_t)(fminf( b.x, 31.999f ) * 2048.0f), bg = (uint32_t)(fminf( b.y, 31.999f ) * 2048.0f), bb = (uint32_t)(fminf( b.z, 31.999f ) * 2048.0f);
	return make_float4( __uint_as_float( (ar << 16) + ag ), __uint_as_float( ab ), __uint_as_float( (br << 16) + bg ), __uint_as_float( bb ) );
}
__device__ __forceinline__ F3 directOf( const float4 v )
{
	const uint32_t a = __float_as_uint( v.x ), b = __float_as_uint( v.y );
	return f3( (float)(a >> 16) * (1.0f / 2048.0f), (float)(a & 65535u) * (1.0f / 2048.0f), (float)b * (1.0f / 2048.0f) );
}
__device__ __forceinline__ F3 indirectOf( const float4 v )
{
	const uint32_t a = __float_as_uint( v.z ), b = __float_as_uint( v.w );
	return f3( (float)(a >> 16) * (1.0f / 2048.0f), (float)(a & 65535u) * (1.0f / 2048.0f), (float)b * (1.0f / 2048.0f) );
}
// :140-160
__device__ __forceinline__ F3 toYCoCg( const F3 c )
{
	const F3 v = minEach( c, 4.0f ); // (the clamp helps AA on strong HDR)
	return f3( dot( v, f3( 1, 2, 1 ) ) * 0.25f, dot( v, f3( 2, 0, -2 ) ) * 0.25f + (0.5f * 256.0f / 255.0f), dot( v, f3( -1, 2, -1 ) ) * 0.25f + (0.5f * 256.0f / 255.0f) );
}
__device__ __forceinline__ F3 fromYCoCg( const F3 c )
{
	const float Y = c.x, Co = c.y - (0.5f * 256.0f / 255.0f), Cg = c.z - (0.5f * 256.0f / 255.0f);
	return f3( Y + Co - Cg, Y + Cg, Y - Co - Cg );
}
__device__ __forceinline__ float luminance( const F3 c ) { return 0.299f * fminf( c.x, 10.0f ) + 0.587f * fminf( c.y, 10.0f ) + 0.114f * fminf( c.z, 10.0f ); }

// ---- history fetches (lib/CUDA/shared_kernel_code/sampling_shared.h:96-205) -------------------------------------------------------
// The four taps of a bilinear history fetch and their weights; a tap counts when its normal is within `minDot` of the pixel's and it
// has the pixel's specularity class (the two flag bits below the packed normal).  Returns the weight sum, 0 = no usable tap.
struct HistoryTaps { int idx[4]; float w[4]; float sum; };
__device__ __forceinline__ bool historyTaps( HistoryTaps& t, const float4* __restrict__ prevWorldPos, const float4 localPos, const F3 localNormal, const float minDot,
	const float u, const float v, const int w, const int h )
{
	const int iu1 = (int)floorf( u ), iv1 = (int)floorf( v ), iu0 = max( 0, iu1 - 1 ), iv0 = max( 0, iv1 - 1 );
	if (iu1 >= w || iv1 >= h || iu1 < 0 || iv1 < 0) return false;
	const float fu = u - floorf( u ), fv = v - floorf( v );
	t.idx[0] = iu0 + iv0 * w, t.idx[1] = iu1 + iv0 * w, t.idx[2] = iu0 + iv1 * w, t.idx[3] = iu1 + iv1 * w;
	t.w[0] = (1 - fu) * (1 - fv), t.w[1] = fu * (1 - fv), t.w[2] = (1 - fu) * fv, t.w[3] = 1 - (t.w[0] + t.w[1] + t.w[2]);
	const uint32_t localClass = __float_as_uint( localPos.w ) & 3u;
#pragma unroll
	for (int k = 0; k < 4; k++)
	{
		const uint32_t n = __float_as_uint( prevWorldPos[t.idx[k]].w );
		if (dot( unpackNormal( n ), localNormal ) < minDot || (n & 3u) != localClass) t.w[k] = 0;
	}
	t.sum = t.w[0] + t.w[1] + t.w[2] + t.w[3];
	return true;
}

// distance between the pixel's world position and the previous frame's at a (fractional) screen position: bilinear over the taps that
// are specular and face the same way (finalize_shared.h:54-84); 1e20 = nothing comparable there
__device__ __forceinline__ float worldDistanceAt( const int px, const int py, const float4 current, const float4* __restrict__ prevWorldPos, const int w, const int h )
{
	const float4 p = (px >= 0 && py >= 0 && px < w && py < h) ? prevWorldPos[px + py * w] : make_float4( 1e20f, 1e20f, 1e20f, __uint_as_float( 0u ) );
	if ((__float_as_uint( p.w ) & 3u) != 1u) return 1e21f;
	if (dot( unpackNormal( __float_as_uint( current.w ) ), unpackNormal( __float_as_uint( p.w ) ) ) < 0.85f) return 1e21f;
	const F3 d = f3( current.x - p.x, current.y - p.y, current.z - p.z );
	return sqrtf( dot( d, d ) );
}
__device__ float fineWorldDistance( const float sx, const float sy, const float4 current, const float4* __restrict__ prevWorldPos, const int w, const int h )
{
	const int x0 = (int)sx, y0 = (int)sy;
	const float fx = sx - floorf( sx ), fy = sy - floorf( sy );
	const float wt[4] = { (1 - fx) * (1 - fy), fx * (1 - fy), (1 - fx) * fy, fx * fy };
	float total = 0, weight = 0;
#pragma unroll
	for (int k = 0; k < 4; k++)
	{
		const float d = worldDistanceAt( x0 + (k & 1), y0 + (k >> 1), current, prevWorldPos, w, h );
		if (d < 1e20f) total += d * wt[k], weight += wt[k];
	}
	return weight == 0 ? 1e20f : total / weight;
}

// ---- prepare ------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__( 256 ) prepareFilterKernel( const FilterFrame f, const FilterReprojection r )
{
	const int x = threadIdx.x + blockIdx.x * blockDim.x, y = threadIdx.y + blockIdx.y * blockDim.y;
	if (x >= f.width || y >= f.height) return;
	const int pixel = x + y * f.width, w = f.width, h = f.height;
	const uint4 feature = f.features[pixel];
	const float4 worldPos = f.worldPos[pixel];
	// demodulate: light = radiance / albedo, clamped, direct and indirect apart
	const F3 albedo = unpackAlbedoMin1( feature.x );
	const F3 reci = f3( 1.0f / albedo.x, 1.0f / albedo.y, 1.0f / albedo.z );
	const F3 direct = minEach( (f3( f.accumulator[pixel] ) * r.pixelValueScale) * reci, r.directClamp );
	const F3 indirect = minEach( (f3( f.accumulator[pixel + w * h] ) * r.pixelValueScale) * reci, r.indirectClamp );
	f.shading[pixel] = packLight( direct, indirect );
	float lumDirect = luminance( direct ), lumDirect2 = lumDirect * lumDirect;
	float lumIndirect = luminance( indirect ), lumIndirect2 = lumIndirect * lumIndirect;
	// where was this surface point in the previous frame?
	float px, py;
	if (((feature.w >> 4) & 3u) == 0)
	{
		// seen directly: through the previous camera
		const F3 toPoint = f3( worldPos.x - r.prevPos.x, worldPos.y - r.prevPos.y, worldPos.z - r.prevPos.z );
		const F3 D = toPoint * rsqrtf( dot( toPoint, toPoint ) );
		const float t = r.prevPos.w / dot( f3( r.prevE ), D );
		const F3 S = f3( r.prevPos.x + D.x * t, r.prevPos.y + D.y * t, r.prevPos.z + D.z * t );
		px = dot( S, f3( r.prevRight ) ) - r.prevRight.w - r.j0, py = dot( S, f3( r.prevUp ) ) - r.prevUp.w - r.j1;
	}
	else
	{
		// seen in a mirror: the same pixel is the best guess for a camera at rest; else a diamond search over the previous world positions
		px = (float)x, py = (float)y;
		if (!r.camIsStationary)
		{
			const float4 prev = f.prevWorldPos[pixel];
			const F3 d = f3( worldPos.x - prev.x, worldPos.y - prev.y, worldPos.z - prev.z );
			float bestDist = sqrtf( dot( d, d ) ), step = 5.0f;
			const float ox = r.j0 - r.prevj0, oy = r.j1 - r.prevj1;
			for (int iter = 0;;)
			{
				// the four neighbours at distance `step`, in the order left, right, up, down; the last one that improves wins
				const float cx[4] = { px - step, px + step, px, px }, cy[4] = { py, py, py - step, py + step };
				int tap = 0;
				float nx = px, ny = py;
#pragma unroll
				for (int k = 0; k < 4; k++)
				{
					const float dist = fineWorldDistance( cx[k] + ox, cy[k] + oy, worldPos, f.prevWorldPos, w, h );
					if (dist < bestDist) bestDist = dist, nx = cx[k], ny = cy[k], tap = k + 1;
				}
				px = nx, py = ny;
				if (tap == 0) { step *= 0.45f; if (step < 0.05f) break; }
				if (++iter == 25) break;
			}
		}
	}
	px += 0.5f, py += 0.5f;
	// luminance moments: 20 % new when the history is usable; the history length lives in the low four bits of features.w
	uint32_t history = feature.w & 15u;
	bool reset = true;
	if (px >= 0 && px < w && py >= 0 && py < h)
	{
		HistoryTaps t;
		if (historyTaps( t, f.prevWorldPos, worldPos, unpackNormal( feature.y ), 0.95f, px, py, w, h ) && t.sum != 0)
		{
			const float4 m0 = f.prevMoments[t.idx[0]], m1 = f.prevMoments[t.idx[1]], m2 = f.prevMoments[t.idx[2]], m3 = f.prevMoments[t.idx[3]];
			const float inv = 1.0f / t.sum;
			const float hx = (t.w[0] * m0.x + t.w[1] * m1.x + t.w[2] * m2.x + t.w[3] * m3.x) * inv, hy = (t.w[0] * m0.y + t.w[1] * m1.y + t.w[2] * m2.y + t.w[3] * m3.y) * inv;
			const float hz = (t.w[0] * m0.z + t.w[1] * m1.z + t.w[2] * m2.z + t.w[3] * m3.z) * inv, hw = (t.w[0] * m0.w + t.w[1] * m1.w + t.w[2] * m2.w + t.w[3] * m3.w) * inv;
			if (hx > -1)
			{
				lumDirect = 0.2f * lumDirect + 0.8f * hx, lumDirect2 = 0.2f * lumDirect2 + 0.8f * hy;
				lumIndirect = 0.2f * lumIndirect + 0.8f * hz, lumIndirect2 = 0.2f * lumIndirect2 + 0.8f * hw;
				if (history < 15) history++;
				reset = false;
			}
		}
	}
	f.features[pixel].w = reset ? (feature.w & 0xfffffff0u) : ((feature.w & 0xfffffff0u) | history);
	f.motion[pixel] = make_float2( px, py );
	f.moments[pixel] = make_float4( lumDirect, lumDirect2, lumIndirect, lumIndirect2 );
}

// ---- reconstruct: one a-trous pass ---------------------------------------------------------------------------------------------------
// A = this pass's input, B = the previous frame's first-pass output (history, phase 1 only), C = output.  5x5 taps without the four
// corners of the outer ring rows (the outer rows are three wide), spread by step = 2^(phase-1).
__global__ void __launch_bounds__( 64, 6 ) reconstructKernel( const FilterFrame f, const float4* __restrict__ A, const float4* __restrict__ B, float4* __restrict__ C,
	const int phase, const int lastPass )
{
	const int x = threadIdx.x + blockIdx.x * blockDim.x, y = threadIdx.y + blockIdx.y * blockDim.y;
	if (x >= f.width || y >= f.height) return;
	const int w = f.width, h = f.height, pixel = x + y * w;
	const uint4 feature = f.features[pixel];
	const F3 localNormal = unpackNormal( feature.y ), localColor = unpackAlbedo( feature.x );
	const uint32_t localMaterial = feature.w >> 4;
	const float4 own = A[pixel];
	F3 directSum = directOf( own ), indirectSum = indirectOf( own );
	float directWeight = 1, indirectWeight = 1;
	const float localDirect = luminance( directSum ), localIndirect = luminance( indirectSum );
	const float localDepth = __uint_as_float( feature.z );
	const float4 dd = f.deltaDepth[pixel];
	const float ddx = dd.z, ddy = dd.w;
	// luminance edge stop from the temporal variance; a pixel without history trusts its neighbours 400 times more
	const float sigma = 10.0f * __uint_as_float( (uint32_t)(127 - (phase - 1)) << 23 );
	const float factor = (feature.w & 15u) == 0 ? 400.0f : 1.0f;
	const float4 m = f.moments[pixel];
	const float varDirect = m.y - m.x * m.x, varIndirect = m.w - m.z * m.z;
	const float kDirect = -1.0f / (sigma * factor * sqrtf( varDirect + 0.00001f ) + 0.00001f);
	const float kIndirect = -1.0f / (sigma * factor * sqrtf( varIndirect + 0.00001f ) + 0.00001f);
	const int step = 1 << (phase - 1);
	for (int vv = -2; vv <= 2; vv++)
	{
		const int v = vv * step + y, reach = abs( vv ) == 2 ? 1 : 2;
		if (v < 0 || v >= h) continue;
		for (int uu = -reach; uu <= reach; uu++) if (uu != 0 || vv != 0)
		{
			const int u = min( max( uu * step + x, 0 ), w - 1 ); // (columns clamp, rows do not: the reference's border behaviour)
			const int tap = u + v * w;
			const float4 light = A[tap];
			const uint4 tapFeature = f.features[tap];
			const float wDist = (float)(uu * uu + vv * vv) * (-1.0f / 7.5f);
			const F3 tapDirect = directOf( light ), tapIndirect = indirectOf( light );
			// normal edge stop.  (powf as in the reference, not seven squarings: on flat regions every tap carries the same quantised light and the
			// weighted mean lands exactly on a step of the 5:11 packing -- a last-bit change of a weight then moves 12 % of the channels by one step.
			// Measured: the squarings save 0.2 of 1.1 ms at 1080p; the bit-equal output was kept.)
			float wNormal = powf( fmaxf( 0.0f, dot( unpackNormal( tapFeature.y ), localNormal ) ), 128 );
			// depth: the error against the plane through the pixel; not too strict, or curved surfaces break up
			const float expected = localDepth + ddx * (float)(uu * step) + ddy * (float)(vv * step);
			const float wDepth = fabsf( expected - __uint_as_float( tapFeature.z ) ) / fmaxf( 0.00001f, (0.5f + phase * 0.5f) * fabsf( expected - localDepth ) );
			// albedo similarity, other materials all but excluded.  (The reference's `? 0.0001 : dot` is a double expression: the product of two
			// floats formed in double and rounded back is the float product, so only the constant's branch needs the double multiply.)
			if ((tapFeature.w >> 4) != localMaterial) wNormal = (float)((double)wNormal * 0.0001);
			else wNormal *= dot( localColor, unpackAlbedo( tapFeature.x ) );
			float wd = wNormal * __expf( fabsf( localDirect - luminance( tapDirect ) ) * kDirect + wDist - wDepth );
			float wi = wNormal * __expf( fabsf( localIndirect - luminance( tapIndirect ) ) * kIndirect + wDist - wDepth );
			if (!isfinite( wd )) wd = 0;
			if (!isfinite( wi )) wi = 0;
			directSum = directSum + tapDirect * wd, directWeight += wd;
			indirectSum = indirectSum + tapIndirect * wi, indirectWeight += wi;
		}
	}
	F3 direct = directSum * (1.0f / fmaxf( 0.0001f, directWeight )), indirect = indirectSum * (1.0f / fmaxf( 0.0001f, indirectWeight ));
	if (phase == 1)
	{
		// temporal: 90 % of the reprojected history, clamped in YCoCg to mean +- 0.75 sigma of the pixel's 3x3 neighbourhood
		const float2 prevPos = f.motion[pixel];
		const int px = (int)prevPos.x, py = (int)prevPos.y;
		HistoryTaps t;
		if (px >= 0 && px < w && py >= 0 && py < h && historyTaps( t, f.prevWorldPos, f.worldPos[pixel], localNormal, 0.975f, prevPos.x, prevPos.y, w, h ) && t.sum != 0)
		{
			const float4 b0 = B[t.idx[0]], b1 = B[t.idx[1]], b2 = B[t.idx[2]], b3 = B[t.idx[3]];
			const float inv = 1.0f / t.sum;
			F3 prevDirect = (t.w[0] * directOf( b0 ) + t.w[1] * directOf( b1 ) + t.w[2] * directOf( b2 ) + t.w[3] * directOf( b3 )) * inv;
			F3 prevIndirect = (t.w[0] * indirectOf( b0 ) + t.w[1] * indirectOf( b1 ) + t.w[2] * indirectOf( b2 ) + t.w[3] * indirectOf( b3 )) * inv;
			if (prevDirect.x != -1)
			{
				prevDirect = toYCoCg( prevDirect ), prevIndirect = toYCoCg( prevIndirect );
				F3 dAvg = toYCoCg( direct ), dVar = dAvg * dAvg, iAvg = toYCoCg( indirect ), iVar = iAvg * iAvg;
				// the window: column -1 needs x > 1, row -1 needs y > 1 (sic), column / row +1 the image edge; always nine in the mean
				for (int k = 0; k < 8; k++)
				{
					// visiting order of the reference: left column (top, middle, bottom), centre column (top, bottom), right column (top, middle, bottom)
					const int ox = k < 3 ? -1 : (k < 5 ? 0 : 1), oy = k < 3 ? k - 1 : (k < 5 ? (k == 3 ? -1 : 1) : k - 6);
					if ((ox < 0 && !(x > 1)) || (ox > 0 && !(x < w - 1)) || (oy < 0 && !(y > 1)) || (oy > 0 && !(y < h - 1))) continue;
					const float4 c4 = A[pixel + ox + oy * w];
					const F3 fd = toYCoCg( directOf( c4 ) ), fi = toYCoCg( indirectOf( c4 ) );
					dAvg = dAvg + fd, dVar = dVar + fd * fd, iAvg = iAvg + fi, iVar = iVar + fi * fi;
				}
				dAvg = dAvg * (1.0f / 9.0f), dVar = dVar * (1.0f / 9.0f), iAvg = iAvg * (1.0f / 9.0f), iVar = iVar * (1.0f / 9.0f);
				const F3 dv = dVar - dAvg * dAvg, iv = iVar - iAvg * iAvg;
				const F3 dSigma = f3( sqrtf( fmaxf( 0.0f, dv.x ) ), sqrtf( fmaxf( 0.0f, dv.y ) ), sqrtf( fmaxf( 0.0f, dv.z ) ) );
				const F3 iSigma = f3( sqrtf( fmaxf( 0.0f, iv.x ) ), sqrtf( fmaxf( 0.0f, iv.y ) ), sqrtf( fmaxf( 0.0f, iv.z ) ) );
				prevDirect = clampEach( prevDirect, dAvg - 0.75f * dSigma, dAvg + 0.75f * dSigma );
				prevIndirect = clampEach( prevIndirect, iAvg - 0.75f * iSigma, iAvg + 0.75f * iSigma );
				direct = direct * 0.1f + fromYCoCg( prevDirect ) * 0.9f;
				indirect = indirect * 0.1f + fromYCoCg( prevIndirect ) * 0.9f;
			}
		}
	}
	if (lastPass)
	{
		// remodulate; the display gamma goes on here because the TAA that follows wants gamma-corrected pixels
		const F3 c = (direct + indirect) * unpackAlbedo( feature.x );
		C[pixel] = make_float4( sqrtf( c.x ), sqrtf( c.y ), sqrtf( c.z ), 1 );
	}
	else C[pixel] = packLight( direct, indirect ); // kept for the next pass and, from phase 1, as the next frame's history
}

// ---- temporal anti-aliasing ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ float mitchellNetravali( const float v ) // B = C = 1/3 (sampling_shared.h:22-27)
{
	const float B = 1.0f / 3.0f, C = 1.0f / 3.0f, x = fabsf( v ), x2 = x * x, x3 = x2 * x;
	if (x < 1) return (1.0f / 6.0f) * ((12 - 9 * B - 6 * C) * x3 + (-18 + 12 * B + 6 * C) * x2 + (6 - 2 * B));
	if (x < 2) return 1.0f / 6.0f * ((-B - 6 * C) * x3 + (6 * B + 30 * C) * x2 + (-12 * B - 48 * C) * x + (8 * B + 24 * C));
	return 0.0f;
}

// pixels: this frame's gamma-corrected image after the last reconstruct pass; out: blended, kept as the next frame's history.
// (The reference blends IN PLACE, finalize_shared.h:383-424: a thread's 3x3 window then holds whichever neighbours other threads have
// already overwritten -- a race whose outcome changes from run to run wherever the variance clamp binds.  Here the window always reads
// the unblended image: the value the reference produces when no neighbour was ahead of the pixel.)
__global__ void __launch_bounds__( 256 ) taaKernel( const float4* __restrict__ pixels, const float4* __restrict__ prevPixels, const float2* __restrict__ motion,
	float4* __restrict__ out, const int w, const int h )
{
	const int x = threadIdx.x + blockIdx.x * blockDim.x, y = threadIdx.y + blockIdx.y * blockDim.y;
	if (x >= w || y >= h) return;
	const int pixel = x + y * w;
	F3 colour = f3( pixels[pixel] );
	const float2 mv = motion[pixel];
	const float u = mv.x - 0.5f, v = mv.y - 0.5f;
	if (u >= 0 && u < w && v >= 0 && v < h)
	{
		// variance clipping (after M. Salvi / C. Wyman) against the 3x3 neighbourhood, in YCoCg
		const F3 fresh = toYCoCg( colour );
		// history through a 4x4 Mitchell-Netravali window (row 0 of the image never contributes: the reference's `y > 0`)
		F3 history;
		{
			const int x1 = (int)(u - 2.0f), y1 = (int)(v - 2.0f);
			float total[4] = { 0, 0, 0, 0 }, totalWeight = 0;
			for (int yy = y1; yy < y1 + 4; yy++) for (int xx = x1; xx < x1 + 4; xx++) if (xx >= 0 && yy > 0 && xx < w && yy < h)
			{
				const float weight = mitchellNetravali( (float)xx - u ) * mitchellNetravali( (float)yy - v );
				const float4 p = prevPixels[xx + yy * w];
				total[0] += p.x * weight, total[1] += p.y * weight, total[2] += p.z * weight, totalWeight += weight;
			}
			const float inv = 1.0f / totalWeight;
			history = toYCoCg( f3( total[0] * inv, total[1] * inv, total[2] * inv ) );
		}
		F3 avg = fresh, var = fresh * fresh;
		for (int k = 0; k < 8; k++) // the window and its order: see reconstructKernel
		{
			const int ox = k < 3 ? -1 : (k < 5 ? 0 : 1), oy = k < 3 ? k - 1 : (k < 5 ? (k == 3 ? -1 : 1) : k - 6);
			if ((ox < 0 && !(x > 1)) || (ox > 0 && !(x < w - 1)) || (oy < 0 && !(y > 1)) || (oy > 0 && !(y < h - 1))) continue;
			const F3 c = toYCoCg( f3( pixels[pixel + ox + oy * w] ) );
			avg = avg + c, var = var + c * c;
		}
		avg = avg * (1.0f / 9.0f), var = var * (1.0f / 9.0f);
		const F3 d = var - avg * avg;
		const F3 sigma = f3( sqrtf( fmaxf( 0.0f, d.x ) ), sqrtf( fmaxf( 0.0f, d.y ) ), sqrtf( fmaxf( 0.0f, d.z ) ) );
		history = clampEach( history, avg - 1.25f * sigma, avg + 1.25f * sigma );
		colour = fromYCoCg( fresh * 0.1f + history * 0.9f );
		if (isnan( colour.x + colour.y + colour.z )) colour = fromYCoCg( fresh );
	}
	out[pixel] = make_float4( fminf( 10.0f, colour.x ), fminf( 10.0f, colour.y ), fminf( 10.0f, colour.z ), 0 );
}

// unsharp mask against the blur of the TAA, then the gamma of the last reconstruct pass is taken off again (the host's post-process
// applies its own).  The one-pixel border of the frame, which the reference leaves as it finds it (finalize_shared.h:438, :475), is black.
__global__ void __launch_bounds__( 256 ) unsharpenKernel( const float4* __restrict__ pixels, float4* __restrict__ frame, const int w, const int h )
{
	const int x = threadIdx.x + blockIdx.x * blockDim.x, y = threadIdx.y + blockIdx.y * blockDim.y;
	if (x >= w || y >= h) return;
	if (x == 0 || y == 0 || x == w - 1 || y == h - 1) { frame[x + y * w] = make_float4( 0, 0, 0, 0 ); return; }
	const float4 p0 = pixels[x - 1 + (y - 1) * w], p1 = pixels[x + (y - 1) * w], p2 = pixels[x + 1 + (y - 1) * w], p3 = pixels[x + 1 + y * w];
	const float4 p4 = pixels[x + 1 + (y + 1) * w], p5 = pixels[x + (y + 1) * w], p6 = pixels[x - 1 + (y + 1) * w], p7 = pixels[x - 1 + y * w];
	const float4 c = pixels[x + y * w];
	float out[3];
	const float cc[3] = { c.x, c.y, c.z };
	const float ring[8][3] = { { p0.x, p0.y, p0.z }, { p1.x, p1.y, p1.z }, { p2.x, p2.y, p2.z }, { p3.x, p3.y, p3.z }, { p4.x, p4.y, p4.z }, { p5.x, p5.y, p5.z }, { p6.x, p6.y, p6.z }, { p7.x, p7.y, p7.z } };
#pragma unroll
	for (int k = 0; k < 3; k++)
	{
		// corners 0.35, edges 0.5, in the order they are listed
		const float blur = 0.35f * ring[0][k] + 0.5f * ring[1][k] + 0.35f * ring[2][k] + 0.5f * ring[3][k] + 0.35f * ring[4][k] + 0.5f * ring[5][k] + 0.35f * ring[6][k] + 0.5f * ring[7][k];
		const float sharp = fmaxf( cc[k], cc[k] * 2.7f - 0.5f * blur );
		out[k] = sharp * sharp;
	}
	frame[x + y * w] = make_float4( out[0], out[1], out[2], 0 );
}

// without TAA.  (The reference takes the square ROOT here, finalize_shared.h:479, of pixels that already carry the gamma of the last
// reconstruct pass; restated as it stands.)
__global__ void __launch_bounds__( 256 ) finishNoTaaKernel( const float4* __restrict__ pixels, float4* __restrict__ frame, const int w, const int h )
{
	const int x = threadIdx.x + blockIdx.x * blockDim.x, y = threadIdx.y + blockIdx.y * blockDim.y;
	if (x >= w || y >= h) return;
	if (x == 0 || y == 0 || x == w - 1 || y == h - 1) { frame[x + y * w] = make_float4( 0, 0, 0, 0 ); return; }
	const float4 p = pixels[x + y * w];
	frame[x + y * w] = make_float4( sqrtf( p.x ), sqrtf( p.y ), sqrtf( p.z ), 0 );
}

dim3 tiles( const int w, const int h, const int bx, const int by ) { return dim3( (w + bx - 1) / bx, (h + by - 1) / by ); }

} // namespace

// ---- launchers ------------------------------------------------------------------------------------------------------------------
// prepareFilter's host part (finalize_shared.h:178-196): the previous camera as a plane + two scaled screen axes
FilterReprojection MakeFilterReprojection( const float* prevView /* pos, p1, p2, p3 */, const int height, const float j0, const float j1, const float prevj0,
	const float prevj1, const uint32_t spp, const float directClamp, const float indirectClamp, const int camIsStationary )
{
	const float *pos = prevView, *p1 = prevView + 3, *p2 = prevView + 6, *p3 = prevView + 9;
	auto sub = []( const float* a, const float* b, float* o ) { o[0] = a[0] - b[0], o[1] = a[1] - b[1], o[2] = a[2] - b[2]; };
	auto dot3 = []( const float* a, const float* b ) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; };
	auto normalise = [&]( float* v ) { const float inv = 1.0f / sqrtf( dot3( v, v ) ); v[0] *= inv, v[1] *= inv, v[2] *= inv; };
	float centre[3] = { 0.5f * (p2[0] + p3[0]), 0.5f * (p2[1] + p3[1]), 0.5f * (p2[2] + p3[2]) }, direction[3], right[3], up[3], span[3];
	sub( centre, pos, direction ), normalise( direction );
	sub( p2, p1, right ), normalise( right );
	sub( p3, p1, up ), normalise( up );
	sub( p3, p1, span );
	const float lenReci = (float)height / sqrtf( dot3( span, span ) );
	FilterReprojection r;
	r.prevPos = make_float4( pos[0], pos[1], pos[2], -(dot3( pos, direction ) - dot3( centre, direction )) );
	r.prevE = make_float4( direction[0], direction[1], direction[2], 0 );
	r.prevRight = make_float4( right[0] * lenReci, right[1] * lenReci, right[2] * lenReci, dot3( p1, right ) * lenReci );
	r.prevUp = make_float4( up[0] * lenReci, up[1] * lenReci, up[2] * lenReci, dot3( p1, up ) * lenReci );
	r.j0 = j0, r.j1 = j1, r.prevj0 = prevj0, r.prevj1 = prevj1;
	r.pixelValueScale = 1.0f / (float)spp, r.directClamp = directClamp, r.indirectClamp = indirectClamp, r.camIsStationary = camIsStationary;
	return r;
}

cudaError_t LaunchFilterPrepare( const FilterFrame& f, const FilterReprojection& r, cudaStream_t s )
{
	prepareFilterKernel<<<tiles( f.width, f.height, 32, 8 ), dim3( 32, 8 ), 0, s>>>( f, r );
	return cudaGetLastError();
}
cudaError_t LaunchFilterReconstruct( const FilterFrame& f, const float4* A, const float4* B, float4* C, int phase, int lastPass, cudaStream_t s )
{
	reconstructKernel<<<tiles( f.width, f.height, 32, 2 ), dim3( 32, 2 ), 0, s>>>( f, A, B, C, phase, lastPass );
	return cudaGetLastError();
}
cudaError_t LaunchFilterTaa( const float4* pixels, const float4* prevPixels, const float2* motion, float4* out, int w, int h, cudaStream_t s )
{
	taaKernel<<<tiles( w, h, 32, 8 ), dim3( 32, 8 ), 0, s>>>( pixels, prevPixels, motion, out, w, h );
	return cudaGetLastError();
}
cudaError_t LaunchFilterUnsharpen( const float4* pixels, float4* frame, int w, int h, cudaStream_t s )
{
	unsharpenKernel<<<tiles( w, h, 32, 8 ), dim3( 32, 8 ), 0, s>>>( pixels, frame, w, h );
	return cudaGetLastError();
}
cudaError_t LaunchFilterFinishNoTaa( const float4* pixels, float4* frame, int w, int h, cudaStream_t s )
{
	finishNoTaaKernel<<<tiles( w, h, 32, 8 ), dim3( 32, 8 ), 0, s>>>( pixels, frame, w, h );
	return cudaGetLastError();
}

} // namespace b200
