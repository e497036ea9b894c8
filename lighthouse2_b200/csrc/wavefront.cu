// wavefront.cu -- the frame around the two waves: ray generation, the shade stage, fused connect, accumulate, finalize, gather.
//
// Re-implements, for the B200 core, the host-driven wave loop of the reference GPU cores
// (lib/rendercore_optix7/rendercore.cpp:621-770; Prime variant lib/RenderCore_OptixPrime_B/rendercore.cpp:463-609):
//   generate primary rays -> [extend -> shade (emits shadow + extension rays, compacted by atomics) -> connect]* -> finalize
// with these differences, all on purpose:
//   * no per-bounce host round trip (the reference copies the counters D->H after every shade, rendercore.cpp:723):
//     every wave reads its ray count from device memory (TraceArgs::countPtr), kernels are persistent grids, so the
//     whole frame is enqueued back to back and the host synchronises once at the end to fill CoreStats
//   * the connect step is fused into the shadow wave (an unoccluded shadow ray adds its potential to the accumulator
//     with atomicAdd; the reference's `accumulator[pixelIdx] += E` is a non-atomic RMW that races for spp > 1,
//     .optix.cu:148, connections.h:33)
//   * output goes to a device RGBA32F buffer (no GL texture exists on this box)
//   * the shade stage follows the reference's shadeKernel (lib/rendercore_optix7/kernels/pathtracer.h:54-240) with either of its
//     BSDFs -- Lambert (lib/sharedBSDFs/lambert.h) or Disney (disney.h, in disney.cuh; Setting "bsdf") --: area / point / spot /
//     directional lights picked by potential contribution, MIS between NEE and emissive hits, postponed BSDF pdf, dielectric
//     transmission (Fresnel, refraction, Beer absorption), interpolated vertex normals for SMOOTH materials, the maps of
//     material_shared.h:100-170 (diffuse trilinear over the MIP chain at the ray-cone LOD with alpha pass-through, detail, normal,
//     detail normal, roughness) from ARGB32 / NRM32 / ARGB128 texels, Russian roulette, HDR skydome lookup, intensity clamp,
//     blue-noise tables supplied by the host (SetBlueNoise), else the WangHash / xor32 stream.  Left out: the specularity map
//     (uploaded but never sampled by the reference either); the consistent-normal correction is applied with the Disney BSDF only
// Pixels can be sharded over several cores/GPUs by interleaved 32x8 tiles (Setting shardIndex/shardCount) or by sample
// (Setting sampleOffset); per-path random streams depend only on (pixel, sample, bounce), so any sharding gives the same
// image and the accumulators of all shards simply add up.
#include "rendercore.h"
#include "../../include/lh2_abi.h"
#include <chrono>
#include <cstring>
#include <vector>

namespace b200 {

namespace {

constexpr int TILE_W = 32, TILE_H = 8, TILE_PIXELS = TILE_W * TILE_H;
constexpr float PI = 3.14159265358979323846f, INVPI = 0.31830988618379067154f, TWOPI = 6.28318530717958647692f;

// Tile numbering.  Tiles are dealt to shards (and, inside a shard, to chains) round robin by tile index.  Numbering the tiles row by
// row would hand every shard the same tile COLUMNS in every row whenever the number of tile columns is a multiple of the shard count
// (3840 / 32 = 120 columns, 8 GPUs): vertical stripes, which follow the scene's content -- the slowest of 8 ranks then rendered 12 % longer
// than the fastest.  So row ty is rotated by ty * skew columns, skew odd and coprime with the shard count: diagonal stripes.
__host__ __device__ __forceinline__ uint32_t tileIndex( const uint32_t tx, const uint32_t ty, const uint32_t tilesX, const uint32_t skew )
{
	return ty * tilesX + (tx + ty * skew) % tilesX;
}
__host__ __device__ __forceinline__ void tileCoords( const uint32_t tile, const uint32_t tilesX, const uint32_t skew, uint32_t& tx, uint32_t& ty )
{
	ty = tile / tilesX;
	tx = (tile % tilesX + tilesX - (ty * skew) % tilesX) % tilesX;
}
static uint32_t tileSkewFor( uint32_t shards )
{
	auto gcd = []( uint32_t a, uint32_t b ) { while (b) { const uint32_t t = a % b; a = b, b = t; } return a; };
	uint32_t s = 3;
	while (gcd( s, shards ? shards : 1 ) != 1) s += 2;
	return s;
}

constexpr int MAX_PATH_LENGTH = 16; // MAXPATHLENGTH of the optix7 core (core_settings.h:24)
// Device-resident wave counters (the reference's Counters, core_settings.h:103-117), uint32 words:
//   [2 len - 1] = extension rays of path length len (written by generate for len = 1, by shade( len - 1 ) otherwise)
//   [2 len]     = shadow rays emitted by shade( len )
// so the two counts one shade launch appends to, [2 len] and [2 len + 1], form ONE aligned 64-bit word and are claimed together.
constexpr int PROBE_BASE = 2 * (MAX_PATH_LENGTH + 1), COUNTER_WORDS = PROBE_BASE + 4; // probe: inst, prim, dist bits, valid
__host__ __device__ __forceinline__ int extCounter( const int len ) { return 2 * len - 1; }
__host__ __device__ __forceinline__ int shadowCounter( const int len ) { return 2 * len; }

// What a frame's kernels need, in two parts.  FrameParams is fixed while nothing but the camera moves: it travels as a kernel parameter
// (constant bank) and is part of the captured frame graph.  PassParams is what changes from frame to frame -- the camera and the
// sample numbering of each pass -- and lives in device memory: the frame (graph or not) starts with one small upload of all its passes.
struct FrameParams {
	float sky[3];                         // constant sky (used when no skydome image is set)
	float epsilon, clampValue;
	uint32_t width, height, pixelCount;
	uint32_t shardIndex, shardCount, tilesX, tilesY, tileSkew; // tile numbering: see tileIndex
	uint32_t chainCount;                  // the shard's tiles are dealt to this many independent chains of waves (local tile t -> chain t % chainCount)
	uint32_t roulette;                    // Russian roulette on (the reference's behaviour) / off (fixed-depth benchmark passes)
	uint32_t jitter, materialCount, bounceCap; // bounceCap: 0 = none, 1 = stop after one diffuse bounce (the reference's ENOUGH_BOUNCES), 2 = two
	int32_t probePixel;
	// lights (lib/RenderSystem/common_classes.h:275-345 as float4 views, lib/CUDA/shared_kernel_code/lights_shared.h)
	const float4* areaLights;  // 6 float4 each
	const float4* pointLights; // 2
	const float4* spotLights;  // 3
	const float4* dirLights;   // 2
	int32_t areaCount, pointCount, spotCount, dirCount;
	const float4* skyPixels;   // skydome image (rgb, 1), skyWidth x skyHeight; null = constant sky
	uint32_t skyWidth, skyHeight;
	const uint32_t* texels;    // ARGB32 texels of all textures, MIP levels behind each other (SetTextures)
	const uint32_t* blueTable; // 5 x 65536 words in the reference's layout (SetBlueNoise); null = none.  A pass uses it when PassParams::blue is set
	// filter mode (Setting "filter"; the reference's lib/RenderCore_Optix7Filter): what the first sample of a pixel finds at its first
	// non-specular vertex, for the reconstruction filter (filter.cu) -- and the accumulator holds two planes, direct and indirect light
	uint4* features;           // x: albedo 10:11:11, y: packed normal + specular flag, z: depth, w: history length | specular << 4 | material << 6; null = filter off
	float4* worldPos;          // xyz + packed normal
	float4* deltaDepth;        // zw: depth derivatives along x and y
};
struct PassParams {
	float pos[3], p1[3], right[3], up[3]; // right = p2 - p1, up = p3 - p1
	float aperture, distortion;           // ViewPyramid::aperture (lens radius), ::distortion (barrel distortion; 0 = none)
	float spreadAngle;                    // ViewPyramid::spreadAngle: ray cone width per unit distance, for the texture LOD
	uint32_t sampleBase, samplesThisPass; // global index of the first sample of this pass
	uint32_t R0;                          // per-frame RNG salt
	uint32_t blue;                        // the blue-noise table applies to this pass (not after the sample index wrapped)
	float finalScale;                     // 1 / samples in the accumulator after this frame (the last pass writes the frame)
	uint32_t blueSlot;                    // filter mode: added to the sample index of the shade stage's blue-noise lookups, + 1 per frame (Optix7Filter/rendercore.cpp:860)
	uint32_t pad[3];
};
static_assert( sizeof( PassParams ) == 96, "PassParams: six float4" );

// One chain of waves: a share of the shard's tiles with its own path buffers, counters and streams.  Chains are independent (disjoint
// pixels), so the tail of one chain's persistent wave kernel is filled by whatever the other chains have in flight.
struct ChainArgs {
	const float4 *rayO, *rayD, *state, *info, *hits;   // this wave's paths
	float4 *rayOout, *rayDout, *stateOut, *infoOut;    // extension rays of the next wave
	float4 *conO, *conD, *conE;                        // shadow rays: E4 = (potential rgb, contribution slot)
	float4 *slotS, *slotC;                             // per-path contribution slots of this pass (see accumulateKernel)
	uint32_t* counters;
	uint32_t chain, chainTiles;                        // chainTiles: local tiles this chain owns
	uint32_t indirectOffset;                           // filter mode: the slots of a path's indirect light (after its diffuse bounce) lie this far behind its direct ones
};

constexpr uint32_t S_SPECULAR = 1, S_BOUNCED = 2, S_VIASPECULAR = 4, S_BOUNCEDTWICE = 8; // pathtracer.h:29-32
constexpr uint32_t PATHIDX_BITS = 28, PATHIDX_MASK = (1u << PATHIDX_BITS) - 1;          // path word = pathIdx | flags << 28

// lib/CUDA/shared_kernel_code/tools_shared.h:60-62
__device__ __forceinline__ uint32_t WangHash( uint32_t s ) { s = (s ^ 61) ^ (s >> 16), s *= 9, s = s ^ (s >> 4), s *= 0x27d4eb2d, s = s ^ (s >> 15); return s; }
__device__ __forceinline__ uint32_t RandomInt( uint32_t& s ) { s ^= s << 13, s ^= s >> 17, s ^= s << 5; return s; }
__device__ __forceinline__ float RandomFloat( uint32_t& s ) { return RandomInt( s ) * 2.3283064365387e-10f; }

// tools_shared.h:332-347 (after E. Heitz et al., "A Low-Discrepancy Sampler that Distributes Monte Carlo Errors as a Blue Noise in
// Screen Space"): table = 256 x 256 Sobol bytes, then scrambling and ranking tiles of 128 x 128 x 8, one word per entry
__device__ __forceinline__ float BlueNoiseSampler( const uint32_t* __restrict__ table, int x, int y, int sampleIndex, int sampleDimension )
{
	x &= 127, y &= 127, sampleIndex &= 255, sampleDimension &= 255;
	const int ranked = (sampleIndex ^ (int)__ldg( table + sampleDimension + (x + y * 128) * 8 + 65536 * 3 )) & 255;
	int value = (int)__ldg( table + sampleDimension + ranked * 256 );
	value ^= (int)__ldg( table + (sampleDimension & 7) + (x + y * 128) * 8 + 65536 );
	return (0.5f + (float)value) * (1.0f / 256.0f);
}

// sin and cos of x >= 0 from separately rounded float operations only (this file is compiled with -fmad=false): quadrant reduction with
// a two-constant pi/2, then degree-9 / degree-8 polynomials on [-pi/4, pi/4]; absolute error about 1e-7.  The reference's cores call the
// hardware approximation __sincosf (tools_shared.h:250-256, .optix.cu:51-64), whose bits no host code can reproduce: with it the frames of
// the CPU restatement (tests/wavefront_ref.py) forked from the device's on a last-bit difference of a bounce direction in a fraction of a
// per cent of the pixels.  tests/wavefront_ref.py::sincos_det restates this function operation by operation, so the two sides agree to the bit.
__device__ __forceinline__ void SinCosDet( const float x, float& s, float& c )
{
	const float kf = rintf( x * 0.636619772f );
	const int k = (int)kf;
	float r = x - kf * 1.5707963705062866f;
	r = r - kf * -4.371139000186243e-8f;
	const float r2 = r * r;
	const float sp = r + (r * r2) * (-0.16666667f + r2 * (0.0083333310f + r2 * (-0.00019840874f + r2 * 2.7525562e-6f)));
	const float cp = 1.0f + r2 * (-0.5f + r2 * (0.041666638f + r2 * (-0.0013888378f + r2 * 2.4760495e-5f)));
	const float ss = (k & 1) ? cp : sp, cc = (k & 1) ? sp : cp;
	s = (k & 2) ? -ss : ss, c = ((k + 1) & 2) ? -cc : cc;
}

__device__ __forceinline__ uint32_t laneMaskLt()
{
	uint32_t m;
	asm( "mov.u32 %0, %%lanemask_lt;" : "=r"( m ) );
	return m;
}

// Block-aggregated appends (stream compaction).  One atomic per CTA and iteration: every append of a wave goes to the same
// address, and same-address atomics serialise in L2 -- with one atomic per WARP the shade kernel spent half its time waiting for
// them (ncu: 47 % of the stall samples on the atomicAdd).  All threads of the block call, every iteration, in uniform control flow.
template <int WARPS> __device__ __forceinline__ uint32_t blockAppend( uint32_t* counter, const bool want, uint32_t* sWarp /* [WARPS + 1] */ )
{
	const uint32_t m = __ballot_sync( 0xffffffffu, want ), warp = threadIdx.x >> 5;
	if ((threadIdx.x & 31) == 0) sWarp[warp] = __popc( m );
	__syncthreads();
	if (threadIdx.x == 0)
	{
		uint32_t total = 0;
		for (int i = 0; i < WARPS; i++) { const uint32_t c = sWarp[i]; sWarp[i] = total, total += c; }
		sWarp[WARPS] = total ? atomicAdd( counter, total ) : 0;
	}
	__syncthreads();
	const uint32_t slot = sWarp[WARPS] + sWarp[warp] + __popc( m & laneMaskLt() );
	__syncthreads(); // sWarp is reused by the next call
	return slot;
}
// two streams at once: counts[0] += #wantA, counts[1] += #wantB with one 64-bit atomic (counts is 8-byte aligned)
template <int WARPS> __device__ __forceinline__ void blockAppend2( uint32_t* counts, const bool wantA, const bool wantB, uint32_t* sWarp /* [2 WARPS + 2] */,
	uint32_t& slotA, uint32_t& slotB )
{
	const uint32_t mA = __ballot_sync( 0xffffffffu, wantA ), mB = __ballot_sync( 0xffffffffu, wantB ), warp = threadIdx.x >> 5;
	if ((threadIdx.x & 31) == 0) sWarp[warp] = __popc( mA ), sWarp[WARPS + warp] = __popc( mB );
	__syncthreads();
	if (threadIdx.x == 0)
	{
		uint32_t tA = 0, tB = 0;
		for (int i = 0; i < WARPS; i++)
		{
			const uint32_t a = sWarp[i], b = sWarp[WARPS + i];
			sWarp[i] = tA, sWarp[WARPS + i] = tB, tA += a, tB += b;
		}
		unsigned long long base = 0;
		if (tA | tB) base = atomicAdd( reinterpret_cast<unsigned long long*>( counts ), (unsigned long long)tA | ((unsigned long long)tB << 32) );
		sWarp[2 * WARPS] = (uint32_t)base, sWarp[2 * WARPS + 1] = (uint32_t)(base >> 32);
	}
	__syncthreads();
	slotA = sWarp[2 * WARPS] + sWarp[warp] + __popc( mA & laneMaskLt() );
	slotB = sWarp[2 * WARPS + 1] + sWarp[WARPS + warp] + __popc( mB & laneMaskLt() );
	__syncthreads();
}

// ---- primary rays (GPU cores: rendercore_optix7/optix/.optix.cu:66-123, RenderCore_OptixPrime_B/kernels/camera.h:40-94) ----------
// u = (x + r0) / w, v = (y + r1) / h, P = p1 + u right + v up, O = point on the lens, D = normalize(P - O); optional barrel distortion.
__global__ void __launch_bounds__( 256 ) generateEyeRaysKernel( const FrameParams f, const PassParams* __restrict__ pass, const ChainArgs c )
{
	__shared__ uint32_t sWarp[9];
	const PassParams P = *pass;
	const uint32_t perSample = c.chainTiles * TILE_PIXELS, total = perSample * P.samplesThisPass;
	const uint32_t stride = gridDim.x * blockDim.x;
	const uint32_t* const blueTable = P.blue ? f.blueTable : nullptr;
	for (uint32_t base = blockIdx.x * blockDim.x; base < total; base += stride) // block-uniform trip count (blockAppend synchronises)
	{
		const uint32_t job = base + threadIdx.x;
		bool valid = job < total;
		uint32_t x = 0, y = 0, sample = 0;
		if (valid)
		{
			sample = job / perSample;
			const uint32_t q = job - sample * perSample, chainTile = q / TILE_PIXELS, inTile = q % TILE_PIXELS;
			const uint32_t tile = (chainTile * f.chainCount + c.chain) * f.shardCount + f.shardIndex;
			uint32_t tx, ty;
			tileCoords( tile, f.tilesX, f.tileSkew, tx, ty );
			x = tx * TILE_W + (inTile % TILE_W), y = ty * TILE_H + inTile / TILE_W;
			valid = x < f.width && y < f.height;
		}
		const uint32_t slot = blockAppend<8>( c.counters + extCounter( 1 ), valid, sWarp );
		if (!valid) continue;
		const uint32_t pixelIdx = x + y * f.width, pathIdx = pixelIdx + (P.sampleBase + sample) * f.pixelCount;
		float r0 = 0.5f, r1 = 0.5f;
		uint32_t seed = WangHash( pathIdx + P.R0 );
		// the first 256 samples of a pixel take their four numbers from the blue-noise table (.optix.cu:66-79)
		const int sampleIdx = (int)(P.sampleBase + sample);
		const bool blue = blueTable && sampleIdx < 256;
		if (f.jitter == 2) r0 = r1 = 0; // filter mode: the pixel's corner; the sub-pixel offset is in p1 (Optix7Filter/optix/.optix.cu:84-88, rendercore.cpp:733-742)
		else if (f.jitter)
		{
			if (blue) r0 = BlueNoiseSampler( blueTable, x, y, sampleIdx, 0 ), r1 = BlueNoiseSampler( blueTable, x, y, sampleIdx, 1 );
			else r0 = RandomFloat( seed ), r1 = RandomFloat( seed );
		}
		// thin lens: a point on a 9-blade aperture of radius `aperture` in the screen-parallel plane through pos (RandomPointOnLens, .optix.cu:51-64)
		float ox = P.pos[0], oy = P.pos[1], oz = P.pos[2];
		if (P.aperture > 0)
		{
			const float l0 = blue ? BlueNoiseSampler( blueTable, x, y, sampleIdx, 2 ) : RandomFloat( seed );
			float l1 = blue ? BlueNoiseSampler( blueTable, x, y, sampleIdx, 3 ) : RandomFloat( seed );
			const float blade = (float)(int)(l0 * 9);
			float l2 = (l0 - blade * (1.0f / 9.0f)) * 9.0f;
			float x1, y1, x2, y2;
			SinCosDet( blade * 3.14159265358979323846f / 4.5f, x1, y1 );
			SinCosDet( (blade + 1.0f) * 3.14159265358979323846f / 4.5f, x2, y2 );
			if (l1 + l2 > 1) l1 = 1.0f - l1, l2 = 1.0f - l2;
			const float xr = x1 * l1 + x2 * l2, yr = y1 * l1 + y2 * l2;
			ox += P.aperture * (P.right[0] * xr + P.up[0] * yr), oy += P.aperture * (P.right[1] * xr + P.up[1] * yr), oz += P.aperture * (P.right[2] * xr + P.up[2] * yr);
		}
		float u, v;
		if (P.distortion == 0) u = ((float)x + r0) / (float)f.width, v = ((float)y + r1) / (float)f.height;
		else
		{
			// barrel distortion of the pixel position about the screen centre (.optix.cu:88-97)
			const float tx = (float)x / (float)f.width - 0.5f, ty = (float)y / (float)f.height - 0.5f;
			const float rr = tx * tx + ty * ty;
			const float rq = sqrtf( rr ) * (1.0f + P.distortion * rr + P.distortion * rr * rr);
			const float theta = atan2f( tx, ty );
			const float bx = (sinf( theta ) * rq + 0.5f) * (float)f.width, by = (cosf( theta ) * rq + 0.5f) * (float)f.height;
			u = (bx + r0) / (float)f.width, v = (by + r1) / (float)f.height;
		}
		const float dx = P.p1[0] + u * P.right[0] + v * P.up[0] - ox;
		const float dy = P.p1[1] + u * P.right[1] + v * P.up[1] - oy;
		const float dz = P.p1[2] + u * P.right[2] + v * P.up[2] - oz;
		const float il = 1.0f / sqrtf( dx * dx + dy * dy + dz * dz );
		c.rayOout[slot] = make_float4( ox, oy, oz, f.epsilon );
		c.rayDout[slot] = make_float4( dx * il, dy * il, dz * il, 1e34f );
		c.stateOut[slot] = make_float4( 1.f, 1.f, 1.f, 1.f );                      // throughput, postponed bsdf pdf
		c.infoOut[slot] = make_float4( __uint_as_float( pathIdx ), 0.f, 0.f, 0.f ); // pathIdx | flags << 28, normal at the previous vertex
	}
}

// lib/CUDA/shared_kernel_code/tools_shared.h:211-220, 250-256, 279-292
__device__ __forceinline__ void Tangent2World( const float vx, const float vy, const float vz, const float nx, const float ny, const float nz,
	float& ox, float& oy, float& oz )
{
	const float sign = copysignf( 1.0f, nz );
	const float a = -1.0f / (sign + nz), b = nx * ny * a;
	const float Bx = 1.0f + sign * nx * nx * a, By = sign * b, Bz = -sign * nx;
	const float Tx = b, Ty = sign + ny * ny * a, Tz = -ny;
	ox = vx * Tx + vy * Bx + vz * nx, oy = vx * Ty + vy * By + vz * ny, oz = vx * Tz + vy * Bz + vz * nz;
}
__device__ __forceinline__ void SafeOrigin( const float ix, const float iy, const float iz, const float rx, const float ry, const float rz,
	const float nx, const float ny, const float nz, const float eps, float& ox, float& oy, float& oz )
{
	const float parallel = 1 - fabsf( nx * rx + ny * ry + nz * rz ), v = parallel * parallel;
	ox = ix + rx * eps * (1 - v) + nx * eps * v, oy = iy + ry * eps * (1 - v) + ny * eps * v, oz = iz + rz * eps * (1 - v) + nz * eps * v;
}

// ---- small vector helpers for the shade stage --------------------------------------------------------------------------------------
struct V3 { float x, y, z; };
__device__ __forceinline__ V3 v3( const float x, const float y, const float z ) { V3 r; r.x = x, r.y = y, r.z = z; return r; }
__device__ __forceinline__ V3 v3( const float4 a ) { return v3( a.x, a.y, a.z ); }
__device__ __forceinline__ V3 operator+( const V3 a, const V3 b ) { return v3( a.x + b.x, a.y + b.y, a.z + b.z ); }
__device__ __forceinline__ V3 operator-( const V3 a, const V3 b ) { return v3( a.x - b.x, a.y - b.y, a.z - b.z ); }
__device__ __forceinline__ V3 operator*( const V3 a, const V3 b ) { return v3( a.x * b.x, a.y * b.y, a.z * b.z ); }
__device__ __forceinline__ V3 operator*( const V3 a, const float s ) { return v3( a.x * s, a.y * s, a.z * s ); }
__device__ __forceinline__ float dot( const V3 a, const V3 b ) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ V3 normalize( const V3 a ) { return a * (1.0f / sqrtf( dot( a, a ) )); }
#include "disney.cuh"

// ---- lights (lib/CUDA/shared_kernel_code/lights_shared.h) ---------------------------------------------------------------------------
// potential contribution of light i to a surface point I with normal N: the importance used to pick a light (:33-94).
// `bary.x < 0` = implicit-connection variant: the light is taken at the given point P (an emissive hit) instead of a sampled point.
__device__ __forceinline__ float PotentialAreaLight( const FrameParams& f, const int i, const V3 I, const V3 N, const V3 P, const V3 bary )
{
	const float4 c4 = __ldg( f.areaLights + i * 6 ), LN = __ldg( f.areaLights + i * 6 + 1 );
	V3 L = P;
	if (bary.x >= 0)
	{
		const float4 V0 = __ldg( f.areaLights + i * 6 + 3 ), V1 = __ldg( f.areaLights + i * 6 + 4 ), V2 = __ldg( f.areaLights + i * 6 + 5 );
		L = v3( V0 ) * bary.x + v3( V1 ) * bary.y + v3( V2 ) * bary.z;
	}
	L = L - I;
	const float att = 1.0f / dot( L, L );
	L = normalize( L );
	return c4.w * fmaxf( 0.0f, -dot( v3( LN ), L ) ) * fmaxf( 0.0f, dot( N, L ) ) * att;
}
__device__ __forceinline__ float PotentialPointLight( const FrameParams& f, const int i, const V3 I, const V3 N )
{
	const float4 p4 = __ldg( f.pointLights + i * 2 );
	const V3 L = v3( p4 ) - I;
	return p4.w * fmaxf( 0.0f, dot( N, L ) ) * (1.0f / dot( L, L )); // L is not normalised here in the reference either (:60-62)
}
__device__ __forceinline__ float PotentialSpotLight( const FrameParams& f, const int i, const V3 I, const V3 N )
{
	const float4 p4 = __ldg( f.spotLights + i * 3 ), r4 = __ldg( f.spotLights + i * 3 + 1 ), d4 = __ldg( f.spotLights + i * 3 + 2 );
	V3 L = v3( p4 ) - I;
	const float att = 1.0f / dot( L, L );
	L = normalize( L );
	const float d = (fmaxf( 0.0f, -dot( L, v3( d4 ) ) ) - r4.w) / (p4.w - r4.w); // p4.w = cos inner, r4.w = cos outer
	return (r4.x + r4.y + r4.z) * fmaxf( 0.0f, fminf( 1.0f, d ) ) * fmaxf( 0.0f, dot( N, L ) ) * att;
}
__device__ __forceinline__ float PotentialDirLight( const FrameParams& f, const int i, const V3 N )
{
	const float4 d4 = __ldg( f.dirLights + i * 2 );
	return d4.w * fmaxf( 0.0f, -dot( v3( d4 ), N ) );
}
// light k of the concatenated list area | point | spot | directional
__device__ __forceinline__ float PotentialLight( const FrameParams& f, int k, const V3 I, const V3 N, const V3 P, const V3 bary )
{
	if (k < f.areaCount) return PotentialAreaLight( f, k, I, N, P, bary );
	k -= f.areaCount;
	if (k < f.pointCount) return PotentialPointLight( f, k, I, N );
	k -= f.pointCount;
	if (k < f.spotCount) return PotentialSpotLight( f, k, I, N );
	return PotentialDirLight( f, k - f.spotCount, N );
}

// :103-120 -- probability that NEE at the previous vertex (O, normal N) would have picked area light `idx`, seen at point I.
// The reference keeps the potentials in a fixed array of 8 (MAXISLIGHTS) and overruns it with more lights; here nothing is stored.
__device__ float LightPickProb( const FrameParams& f, const int idx, const V3 O, const V3 N, const V3 I )
{
	const int lights = f.areaCount + f.pointCount + f.spotCount + f.dirCount;
	const V3 none = v3( -1.f, -1.f, -1.f );
	float sum = 0, mine = 0;
	for (int k = 0; k < lights; k++)
	{
		const float c = PotentialLight( f, k, O, N, I, none );
		sum += c;
		if (k == idx) mine = c;
	}
	return sum <= 0 ? 0.f : mine / sum;
}

// :122-142 -- uniform point on a triangle from ONE random number (base-4 digits walk a 4-way subdivision)
__device__ V3 RandomBarycentrics( const float r0 )
{
	const uint32_t uf = (uint32_t)(r0 * 4294967296.0f);
	float ax = 1, ay = 0, bx = 0, by = 1, cx = 0, cy = 0;
	// every digit picks one of four sub-triangles; written with selects, not branches: the four cases are mixed within a warp (the
	// branchy form ran all four per step and was a third of the shade kernel's instructions).  Same sums, same bits.
#pragma unroll 4
	for (int i = 0; i < 16; i++)
	{
		const int d = (uf >> (2 * (15 - i))) & 3;
		const float abx = (ax + bx) * 0.5f, aby = (ay + by) * 0.5f, acx = (ax + cx) * 0.5f, acy = (ay + cy) * 0.5f;
		const float bcx = (bx + cx) * 0.5f, bcy = (by + cy) * 0.5f;
		const float nax = d == 0 ? bcx : d == 1 ? ax : d == 2 ? abx : acx, nay = d == 0 ? bcy : d == 1 ? ay : d == 2 ? aby : acy;
		const float nbx = d == 0 ? acx : d == 1 ? abx : d == 2 ? bx : bcx, nby = d == 0 ? acy : d == 1 ? aby : d == 2 ? by : bcy;
		const float ncx = d == 0 ? abx : d == 1 ? acx : d == 2 ? bcx : cx, ncy = d == 0 ? aby : d == 1 ? acy : d == 2 ? bcy : cy;
		ax = nax, ay = nay, bx = nbx, by = nby, cx = ncx, cy = ncy;
	}
	const float rx = (ax + bx + cx) * 0.3333333f, ry = (ay + by + cy) * 0.3333333f;
	return v3( rx, ry, 1 - rx - ry );
}

// :144-234 -- pick a light by potential contribution, return the sampled point on it
__device__ V3 RandomPointOnLight( const FrameParams& f, const float r0, float r1, const V3 I, const V3 N, float& pickProb, float& lightPdf, V3& lightColor )
{
	const int lights = f.areaCount + f.pointCount + f.spotCount + f.dirCount;
	const V3 bary = RandomBarycentrics( r0 ), zero = v3( 0.f, 0.f, 0.f );
	lightPdf = 0, pickProb = 0, lightColor = zero;
	float sum = 0;
	for (int k = 0; k < lights; k++) sum += PotentialLight( f, k, I, N, zero, bary );
	if (!(sum > 0)) return v3( 1.f, 1.f, 1.f );
	r1 *= sum;
	int k = 0;
	float total = 0, chosen = 0;
	bool found = false;
	for (int i = 0; i < lights; i++)
	{
		const float c = PotentialLight( f, i, I, N, zero, bary );
		total += c;
		if (i == 0) chosen = c;
		if (!found && total >= r1) k = i, chosen = c, found = true;
	}
	pickProb = chosen / sum;
	if (k < f.areaCount)
	{
		const float4 LN = __ldg( f.areaLights + k * 6 + 1 ), rad = __ldg( f.areaLights + k * 6 + 2 );
		const float4 V0 = __ldg( f.areaLights + k * 6 + 3 ), V1 = __ldg( f.areaLights + k * 6 + 4 ), V2 = __ldg( f.areaLights + k * 6 + 5 );
		lightColor = v3( rad );
		const V3 P = v3( V0 ) * bary.x + v3( V1 ) * bary.y + v3( V2 ) * bary.z;
		V3 L = I - P; // from the light to the surface point
		const float sqDist = dot( L, L );
		L = normalize( L );
		const float LNdotL = dot( L, v3( LN ) );
		lightPdf = (LNdotL > 0 && dot( L, N ) < 0) ? sqDist / (LN.w * LNdotL) : 0; // LN.w = area
		return P;
	}
	k -= f.areaCount;
	if (k < f.pointCount)
	{
		const float4 p4 = __ldg( f.pointLights + k * 2 ), rad = __ldg( f.pointLights + k * 2 + 1 );
		lightColor = v3( rad ); // the reference assigns a shadowing local here (:196) and leaves the output unset; the radiance is what is meant
		const V3 L = I - v3( p4 );
		lightPdf = dot( L, N ) < 0 ? dot( L, L ) : 0;
		return v3( p4 );
	}
	k -= f.pointCount;
	if (k < f.spotCount)
	{
		const float4 p4 = __ldg( f.spotLights + k * 3 ), r4 = __ldg( f.spotLights + k * 3 + 1 ), d4 = __ldg( f.spotLights + k * 3 + 2 );
		V3 L = I - v3( p4 );
		const float sqDist = dot( L, L );
		L = normalize( L );
		const float d = (fmaxf( 0.0f, dot( L, v3( d4 ) ) ) - r4.w) / (p4.w - r4.w);
		const float LNdotL = fminf( 1.0f, d );
		lightPdf = (LNdotL > 0 && dot( L, N ) < 0) ? sqDist / LNdotL : 0;
		lightColor = v3( r4 );
		return v3( p4 );
	}
	k -= f.spotCount;
	const float4 d4 = __ldg( f.dirLights + k * 2 ), rad = __ldg( f.dirLights + k * 2 + 1 );
	lightColor = v3( rad );
	lightPdf = dot( v3( d4 ), N ) < 0 ? 1.f : 0.f;
	return I - v3( d4 ) * 1000.0f;
}

// tools_shared.h:185-192 (formulas by Paul Debevec): nearest texel of the equirectangular skydome
__device__ __forceinline__ V3 SampleSkydome( const FrameParams& f, const V3 D )
{
	if (!f.skyPixels) return v3( f.sky[0], f.sky[1], f.sky[2] );
	const uint32_t u = (uint32_t)((float)f.skyWidth * 0.5f * (1.0f + atan2f( D.x, -D.z ) * INVPI));
	const uint32_t v = (uint32_t)((float)f.skyHeight * acosf( D.y ) * INVPI);
	const uint32_t idx = u + v * f.skyWidth;
	return idx < f.skyWidth * f.skyHeight ? v3( __ldg( f.skyPixels + idx ) ) : v3( 0.f, 0.f, 0.f );
}

// lib/CUDA/shared_kernel_code/sampling_shared.h:36-76 (FetchTexel, BILINEAR as the optix7 core builds it) for ARGB32 storage
__device__ __forceinline__ float4 TexelToFloat4( const uint32_t t )
{
	const float r = 1.0f / 256.0f;
	return make_float4( (float)(t & 255u) * r, (float)((t >> 8) & 255u) * r, (float)((t >> 16) & 255u) * r, (float)(t >> 24) * r );
}
// The texel pool.  ARGB32 / NRM32 textures sit in it as one word per texel; an ARGB128 (HDR) texture as four floats per texel
// at a 16-byte aligned place, its offset counted in float4s and marked by bit 31 (SetTextures).  The reference keeps three pools
// (rendercore.cpp:434-505) and its FetchTexel takes the storage as an argument (sampling_shared.h:35-36) -- which none of its callers
// passes as ARGB128 (material_shared.h:121-169, sampling_shared.h:218-259), so there a material that names an HDR texture reads the
// argb32 pool at the HDR texture's offset; here it gets the HDR texels.
constexpr uint32_t TEXEL_HDR = 0x80000000u;
__device__ float4 FetchTexel( const uint32_t* __restrict__ texels, const float tu, const float tv, const uint32_t o, const int w, const int h )
{
	const float tcx = fmaxf( tu + 1000, 0.0f ) * (float)w - 0.5f, tcy = fmaxf( tv + 1000, 0.0f ) * (float)h - 0.5f;
	const int iu = ((int)tcx) % w, iv = ((int)tcy) % h;
	const float fu = tcx - floorf( tcx ), fv = tcy - floorf( tcy );
	const float w0 = (1 - fu) * (1 - fv), w1 = fu * (1 - fv), w2 = (1 - fu) * fv, w3 = 1 - (w0 + w1 + w2);
	const int iu1 = (iu + 1) % w, iv1 = (iv + 1) % h;
	const int i0 = iu + iv * w, i1 = iu1 + iv * w, i2 = iu + iv1 * w, i3 = iu1 + iv1 * w;
	float4 p0, p1, p2, p3;
	if (o & TEXEL_HDR)
	{
		const float4* __restrict__ hdr = reinterpret_cast<const float4*>( texels ) + (o & ~TEXEL_HDR);
		p0 = __ldg( hdr + i0 ), p1 = __ldg( hdr + i1 ), p2 = __ldg( hdr + i2 ), p3 = __ldg( hdr + i3 );
	}
	else
	{
		const uint32_t* __restrict__ ldr = texels + o;
		p0 = TexelToFloat4( __ldg( ldr + i0 ) ), p1 = TexelToFloat4( __ldg( ldr + i1 ) ), p2 = TexelToFloat4( __ldg( ldr + i2 ) ), p3 = TexelToFloat4( __ldg( ldr + i3 ) );
	}
	return make_float4( p0.x * w0 + p1.x * w1 + p2.x * w2 + p3.x * w3, p0.y * w0 + p1.y * w1 + p2.y * w2 + p3.y * w3,
		p0.z * w0 + p1.z * w1 + p2.z * w2 + p3.z * w3, p0.w * w0 + p1.w * w1 + p2.w * w2 + p3.w * w3 );
}
// :78-93 (FetchTexelTrilinear); `levels` is the texture's own MIP count where the reference assumes MIPLEVELCOUNT = 5
__device__ float4 FetchTexelTrilinear( const uint32_t* __restrict__ texels, const float lambda, const float tu, const float tv, const uint32_t offset,
	const int width, const int height, const int levels )
{
	const int level0 = min( levels - 1, (int)lambda ), level1 = min( levels - 1, level0 + 1 );
	const float fl = lambda - floorf( lambda );
	uint32_t o0 = offset, o1 = offset;
	int w0 = width, h0 = height, w1 = width, h1 = height;
	for (int i = 0; i < level0; i++) o0 += w0 * h0, w0 >>= 1, h0 >>= 1;
	for (int i = 0; i < level1; i++) o1 += w1 * h1, w1 >>= 1, h1 >>= 1;
	const float4 p0 = FetchTexel( texels, tu, tv, o0, w0, h0 ), p1 = FetchTexel( texels, tu, tv, o1, w1, h1 );
	return make_float4( (1 - fl) * p0.x + fl * p1.x, (1 - fl) * p0.y + fl * p1.y, (1 - fl) * p0.z + fl * p1.z, (1 - fl) * p0.w + fl * p1.w );
}

// material flag bits of the device record: SMOOTH, HASALPHA (host_material.h:66-71), then one bit per map that is present
constexpr uint32_t M_DIFFUSEMAP = 4, M_DIFFUSEMAP2 = 8, M_NORMALMAP = 16, M_NORMALMAP2 = 32, M_ROUGHNESSMAP = 64;
constexpr uint32_t M_ANYMAP = M_DIFFUSEMAP | M_DIFFUSEMAP2 | M_NORMALMAP | M_NORMALMAP2 | M_ROUGHNESSMAP;
// one extra map of a material: rec[0] = (uv scale, uv offset), rec[1] = (first texel, width, height, xy scale of a normal map)
__device__ __forceinline__ float4 FetchMap( const uint32_t* __restrict__ texels, const float4* __restrict__ rec, const float tu, const float tv )
{
	const float4 uv = __ldg( rec ), d = __ldg( rec + 1 );
	return FetchTexel( texels, uv.x * (uv.z + tu), uv.y * (uv.w + tv), __float_as_uint( d.x ), (int)__float_as_uint( d.y ), (int)__float_as_uint( d.z ) );
}

// core_settings.h:190-192 (CLAMPINTENSITY) and FIXNAN_FLOAT3
__device__ __forceinline__ V3 ClampAndFix( V3 c, const float clampValue )
{
	const float m = fmaxf( c.x, fmaxf( c.y, c.z ) );
	if (m > clampValue) c = c * (clampValue / m);
	if (!isfinite( c.x + c.y + c.z )) c = v3( 0.f, 0.f, 0.f );
	return c;
}

// ---- shade (lib/rendercore_optix7/kernels/pathtracer.h:54-240, BSDF = lib/sharedBSDFs/lambert.h without transmission) -------------
// Path state: rayO/rayD (the extension ray), pathState = (throughput rgb, postponed bsdf pdf), pathInfo = (pathIdx | flags << 28,
// normal at the previous vertex).  One thread per path; new extension / shadow rays are compacted with warp-aggregated appends.
#ifndef B200_SHADE_MIN_BLOCKS
#define B200_SHADE_MIN_BLOCKS 5
#endif
#ifndef B200_SHADE_COMPACT
#define B200_SHADE_COMPACT 0 // 1: the paths of a block are shaded hits first, misses after (see shadeKernel).  Measured on the 4K x 16 spp frame: lanes per
                             // instruction go up, time does not (shade 104.1 -> 105.1 ms summed over the chains, frame 126.6 -> 127.4 ms): the kernel
                             // waits for its scattered record loads, not for issue slots.  Off.
#endif
#ifndef B200_SHADE_GRID
#define B200_SHADE_GRID 2 // resident sets of CTAs per launch
#endif
// filter-mode helpers: packing formats of the feature buffers (lib/CUDA/shared_kernel_code/tools_shared.h:122-183)
__device__ __forceinline__ uint32_t PackNormal2( const V3 N )
{
	const uint32_t x = min( (uint32_t)((N.x + 1) * 511), 1023u ), y = min( (uint32_t)((N.y + 1) * 511), 1023u ), z = min( (uint32_t)((N.z + 1) * 511), 1023u );
	return (x << 2) + (y << 12) + (z << 22);
}
__device__ __forceinline__ uint32_t HDRtoRGB32( const V3 c )
{
	return ((uint32_t)(1023.0f * fminf( 1.0f, c.x )) << 22) + ((uint32_t)(2047.0f * fminf( 1.0f, c.y )) << 11) + (uint32_t)(2047.0f * fminf( 1.0f, c.z ));
}
__device__ __forceinline__ V3 RGB32toHDR( const uint32_t c )
{
	return v3( (float)(c >> 22) * (1.0f / 1023.0f), (float)((c >> 11) & 2047u) * (1.0f / 2047.0f), (float)(c & 2047u) * (1.0f / 2047.0f) );
}

// FILTER: the shade stage of the reference's filtering core (lib/RenderCore_Optix7Filter/kernels/pathtracer.h): the same path tracer, plus
// the feature buffers written by the first sample of every pixel and light split into what arrives before and after the diffuse bounce
template <bool DISNEY, bool FILTER> __global__ void __launch_bounds__( 128, DISNEY ? 4 : B200_SHADE_MIN_BLOCKS ) shadeKernel( const FrameParams f, const PassParams* __restrict__ pass, const int pathLength, const int maxPathLength,
	const ChainArgs c, const float4* __restrict__ shade, const float4* __restrict__ vertexNormals, const float4* __restrict__ instShade, const float4* __restrict__ materials )
{
	__shared__ uint32_t sWarp[2 * 4 + 2];
#if B200_SHADE_COMPACT
	__shared__ uint32_t sGroup[8];
	__shared__ uint8_t sOrder[128];
#endif
	uint32_t* const counters = c.counters;
	const uint32_t count = counters[extCounter( pathLength )];
	// (PassParams fields are read where they are used: uniform, cached loads that cost less than registers held across the whole loop body)
	const uint32_t stride = gridDim.x * blockDim.x;
	for (uint32_t base = blockIdx.x * blockDim.x; base < count; base += stride) // block-uniform trip count (blockAppend2 synchronises)
	{
#if B200_SHADE_COMPACT
		// The block's 128 paths, regrouped: those that hit something first, those that left the scene behind them.  On the bounce waves half
		// of the paths end on the sky after a handful of instructions while their neighbours in the warp work through materials, next event
		// estimation and the BSDF (13-19 of 32 lanes per instruction before this); grouped, the long path runs in full warps.  The order
		// in which paths are shaded changes nothing in the image: every path adds to slots of its own (accumulateKernel).
		uint32_t job;
		{
			const uint32_t mine = base + threadIdx.x;
			const bool valid = mine < count, isHit = valid && (int)__float_as_uint( __ldg( &c.hits[mine].z ) ) >= 0;
			const uint32_t mh = __ballot_sync( 0xffffffffu, isHit ), mm = __ballot_sync( 0xffffffffu, valid && !isHit ), warp = threadIdx.x >> 5;
			if ((threadIdx.x & 31) == 0) sGroup[warp] = __popc( mh ), sGroup[4 + warp] = __popc( mm );
			__syncthreads();
			uint32_t hitsBefore = 0, missesBefore = 0, hitsTotal = 0;
#pragma unroll
			for (uint32_t i = 0; i < 4; i++)
			{
				if (i < warp) hitsBefore += sGroup[i], missesBefore += sGroup[4 + i];
				hitsTotal += sGroup[i];
			}
			if (isHit) sOrder[hitsBefore + __popc( mh & laneMaskLt() )] = (uint8_t)threadIdx.x;
			else if (valid) sOrder[hitsTotal + missesBefore + __popc( mm & laneMaskLt() )] = (uint8_t)threadIdx.x;
			__syncthreads();
			// (threads beyond the wave's last path keep an index past the end, as before)
			job = base + threadIdx.x < count ? base + sOrder[threadIdx.x] : base + threadIdx.x;
		}
#else
		const uint32_t job = base + threadIdx.x;
#endif
		bool wantShadow = false, wantExt = false;
		V3 so = v3( 0.f, 0.f, 0.f ), sl = so, se = so, eo = so, er = so, et = so, en = so;
		float sdist = 0, epdf = 0;
		uint32_t pathIdx = 0, pixelIdx = 0, flags = 0, slot = 0;
		if (job < count)
		{
			const float4 O4 = __ldg( c.rayO + job ), D4 = __ldg( c.rayD + job ), T4 = __ldg( c.state + job ), P4 = __ldg( c.info + job ), hit = __ldg( c.hits + job );
			pathIdx = __float_as_uint( P4.x ) & PATHIDX_MASK, flags = __float_as_uint( P4.x ) >> PATHIDX_BITS;
			const uint32_t sampleIdxU = pathIdx / f.pixelCount;
			pixelIdx = pathIdx - sampleIdxU * f.pixelCount;
			{
				// contribution slot of this path = its index when the pass was generated (generateEyeRaysKernel's `job`)
				const uint32_t pxX = pixelIdx % f.width, pxY = pixelIdx / f.width;
				const uint32_t tile = tileIndex( pxX / TILE_W, pxY / TILE_H, f.tilesX, f.tileSkew ), chainTile = tile / f.shardCount / f.chainCount;
				slot = (sampleIdxU - __ldg( &pass->sampleBase )) * (c.chainTiles * TILE_PIXELS) + chainTile * TILE_PIXELS + (pxY % TILE_H) * TILE_W + pxX % TILE_W;
			}
			const int prim = (int)__float_as_uint( hit.z );
			const V3 D = v3( D4 ), O = v3( O4 ), lastN = v3( P4.y, P4.z, P4.w );
			V3 throughput = v3( T4 );
			const float bsdfPdf = T4.w;
			// The path's own contribution slots (see accumulateKernel).  The first vertex initialises both by plain stores (every path
			// passes here exactly once, before its first connect wave): slotS with what the vertex adds itself (skydome, a light that was
			// hit) and the depth for the DOF / filter passes (pathtracer.h:74), slotC with zero.  Later vertices add.  One kernel per wave
			// touches a slot at most once and the waves are ordered: no race, fixed order.
			const float depth = prim < 0 ? 10000.f : hit.w;
			bool stored = false;
			// filter mode: what a path gathers after its diffuse bounce is indirect light and goes to the second plane (pathtracer.h:93)
			if (FILTER && (flags & S_BOUNCED)) slot += c.indirectOffset;
			auto contribute = [&]( const V3 v ) {
				if (pathLength == 1) c.slotS[slot] = make_float4( v.x, v.y, v.z, depth ), stored = true;
				else { float* px = reinterpret_cast<float*>( c.slotS + slot ); atomicAdd( px, v.x ), atomicAdd( px + 1, v.y ), atomicAdd( px + 2, v.z ); }
			};
			if (pathLength == 1) c.slotC[slot] = make_float4( 0.f, 0.f, 0.f, 0.f );
			if (FILTER && pathLength == 1) c.slotS[slot + c.indirectOffset] = c.slotC[slot + c.indirectOffset] = make_float4( 0.f, 0.f, 0.f, 0.f );
			// filter mode: the feature record of the pixel, written by its first sample up to the first non-specular vertex (:96-104)
			const bool firstHit = FILTER && !(flags & S_BOUNCED) && sampleIdxU == 0;
			const uint32_t viaSpecular = (flags & S_VIASPECULAR) ? 1u : 0u;
			auto packFeatures = [&]( const uint32_t albedo, const uint32_t packedNormal, const uint32_t material ) {
				uint4* ft = f.features + pixelIdx;
				*ft = make_uint4( albedo, packedNormal, __float_as_uint( hit.w ), (viaSpecular << 4) + (material << 6) + (ft->w & 15u) ); // (the history length stays)
			};
			auto depthDerivatives = [&]( const V3 P, const V3 n ) {
				// the depth the neighbouring pixels' rays would find on the plane of this surface (:48-58; plane and camera both in world space here --
				// the reference takes the triangle from object space, the same thing for untransformed instances)
				const V3 pos = v3( __ldg( &pass->pos[0] ), __ldg( &pass->pos[1] ), __ldg( &pass->pos[2] ) ), p1 = v3( __ldg( &pass->p1[0] ), __ldg( &pass->p1[1] ), __ldg( &pass->p1[2] ) );
				const V3 right = v3( __ldg( &pass->right[0] ), __ldg( &pass->right[1] ), __ldg( &pass->right[2] ) ), up = v3( __ldg( &pass->up[0] ), __ldg( &pass->up[1] ), __ldg( &pass->up[2] ) );
				const float px = (float)(pixelIdx % f.width), py = (float)(pixelIdx / f.width), plane = dot( P - pos, n );
				const V3 dx = normalize( (p1 + right * ((px + 0.5f + 1) * (1.0f / (float)f.width)) + up * ((py + 0.5f) * (1.0f / (float)f.height))) - pos );
				const V3 dy = normalize( (p1 + right * ((px + 0.5f) * (1.0f / (float)f.width)) + up * ((py + 0.5f + 1) * (1.0f / (float)f.height))) - pos );
				f.deltaDepth[pixelIdx] = make_float4( 0, 0, plane / dot( n, dx ) - hit.w, plane / dot( n, dy ) - hit.w );
			};
			if (firstHit && pathLength == 1)
			{
				f.features[pixelIdx] = make_uint4( 0, 0, __float_as_uint( 1e34f ), f.features[pixelIdx].w & 15u );
				f.worldPos[pixelIdx] = make_float4( 0, 0, 0, 0 ), f.deltaDepth[pixelIdx] = make_float4( 0, 0, 0, 0 );
			}
			if (prim < 0)
			{
				// the path left the scene: skydome (pathtracer.h:77-84)
				const V3 sky = ClampAndFix( throughput * SampleSkydome( f, D ) * (1.0f / bsdfPdf), f.clampValue );
				contribute( sky );
				if (firstHit)
				{
					// the path ends before a diffuse vertex: something reasonable for the filter (:115-123)
					const uint32_t pn = PackNormal2( D * -1.0f ) + viaSpecular;
					packFeatures( HDRtoRGB32( sky ), pn, 0 );
					const V3 far = O + D * 50000.0f;
					f.worldPos[pixelIdx] = make_float4( far.x, far.y, far.z, __uint_as_float( pn ) ), f.deltaDepth[pixelIdx] = make_float4( 0, 0, 0, 0 );
				}
			}
			else
			{
				const uint32_t inst = __float_as_uint( hit.y );
				if (pathLength == 1 && (int)pixelIdx == f.probePixel && pathIdx < f.pixelCount)
					counters[PROBE_BASE] = inst, counters[PROBE_BASE + 1] = (uint32_t)prim, counters[PROBE_BASE + 2] = __float_as_uint( hit.w ), counters[PROBE_BASE + 3] = 1; // pathtracer.h:94-97
				const float4 m0 = __ldg( instShade + inst * 4 ), m1 = __ldg( instShade + inst * 4 + 1 ), m2 = __ldg( instShade + inst * 4 + 2 ),
							 info = __ldg( instShade + inst * 4 + 3 );
				const float4 sh = __ldg( shade + __float_as_uint( info.x ) + prim );
				// world-space geometric normal: normalize( transform.TransformVector( N ) ) (bvh.cpp:616-619); also stands in for the
				// interpolated and the normal-mapped normal (iN, fN) of GetShadingData
				const V3 N = normalize( v3( m0.x * sh.x + m0.y * sh.y + m0.z * sh.z, m1.x * sh.x + m1.y * sh.y + m1.z * sh.z, m2.x * sh.x + m2.y * sh.y + m2.z * sh.z ) );
				const uint32_t mat = __float_as_uint( sh.w );
				// material table, MAT_STRIDE float4 per material: (albedo rgb, roughness) (transmission, eta, flag bits, MIP levels)
				// (extinction rgb, first texel) (uv scale, uv offset) (texture width, height, -, -) (Disney parameter words x y z, -);
				// flags: 1 SMOOTH, 2 HASALPHA, 4 diffuse map
				V3 color = v3( 0.8f, 0.8f, 0.8f ), absorption = v3( 0.f, 0.f, 0.f );
				float roughness = 1.0f, transmission = 0.0f, eta = 1.0f;
				uint32_t mflags = 0;
				float4 c1 = make_float4( 0, 0, 0, 0 ), c2 = c1, c3 = c1, c4 = c1;
				disney::Data dd;
				if (DISNEY) disney::Unpack( dd, 0x80800000u /* specular 0.5, roughness 0.5 */, 0u, 0u ), dd.eta = 1.0f;
				if (mat < f.materialCount)
				{
					const float4 c0 = __ldg( materials + mat * MAT_STRIDE );
					c1 = __ldg( materials + mat * MAT_STRIDE + 1 ), c2 = __ldg( materials + mat * MAT_STRIDE + 2 );
					color = v3( c0 ), roughness = c0.w, transmission = c1.x, eta = c1.y, mflags = __float_as_uint( c1.z ), absorption = v3( c2 );
					if (DISNEY)
					{
						const float4 c5 = __ldg( materials + mat * MAT_STRIDE + 5 );
						disney::Unpack( dd, __float_as_uint( c5.x ), __float_as_uint( c5.y ), __float_as_uint( c5.z ) );
						dd.eta = eta, roughness = dd.roughness, transmission = dd.transmission; // the 0.8 fixed-point values drive the specular test too
					}
				}
				if (DISNEY) disney::Tint( dd, color );
				// per-triangle attributes (CoreTri): vertex normals, texture coordinates, LOD -- 4 float4: (vN0,u0) (vN1,u1) (vN2,u2) (v0,v1,v2,LOD)
				const uint32_t bary = __float_as_uint( hit.x );
				const float bu = (float)(bary & 0xffffu) * (1.0f / 65535.0f), bv = (float)(bary >> 16) * (1.0f / 65535.0f), bw = 1 - (bu + bv);
				float4 a0 = make_float4( 0, 0, 0, 0 ), a1 = a0, a2 = a0, a3 = a0, a4 = a0, a5 = a0;
				float alpha2 = 0;
				if (((mflags & (1u | M_ANYMAP)) || DISNEY) && vertexNormals)
				{
					const float4* ta = vertexNormals + (size_t)(__float_as_uint( info.x ) + prim) * TRI_ATTR_STRIDE;
					a0 = __ldg( ta ), a1 = __ldg( ta + 1 ), a2 = __ldg( ta + 2 ), a3 = __ldg( ta + 3 );
					if (DISNEY || (mflags & M_NORMALMAP)) a4 = __ldg( ta + 4 ), a5 = __ldg( ta + 5 ); // (tangent, alpha0) (bitangent, alpha1)
					if (DISNEY) alpha2 = __ldg( ta + 6 ).x;
				}
				// interpolated normal (GetShadingData, material_shared.h:75-81: w N0 + u N1 + v N2 for SMOOTH materials), transformed like N;
				// it also stands in for the normal-mapped normal fN.  (Consistent normals, :90-97, need the per-vertex alpha of the
				// host mesh and are not applied.)
				V3 sN = N;
				if (mflags & 1u)
				{
					const V3 on = v3( a0 ) * bw + v3( a1 ) * bu + v3( a2 ) * bv;
					if (dot( on, on ) > 0)
						sN = normalize( v3( m0.x * on.x + m0.y * on.y + m0.z * on.z, m1.x * on.x + m1.y * on.y + m1.z * on.z, m2.x * on.x + m2.y * on.y + m2.z * on.z ) );
				}
				V3 tangent = v3( a4 );
				if (DISNEY)
				{
					// consistent normal interpolation (material_shared.h:87-97) with the per-vertex alphas of the host mesh
					if (vertexNormals)
					{
						const bool backSide = dot( D, N ) > 0;
						const float alpha = bw * a4.w + bu * a5.w + bv * alpha2;
						sN = disney::ConsistentNormal( D * -1.0f, backSide ? sN * -1.0f : sN, alpha ) * (backSide ? -1.0f : 1.0f);
					}
					// the reference takes CoreTri::T as it is (object space); a mesh without CoreTris gets an arbitrary tangent
					if (!(dot( tangent, tangent ) > 0)) tangent = fabsf( sN.x ) > 0.9f ? v3( 0.f, 1.f, 0.f ) : v3( 1.f, 0.f, 0.f );
				}
				const V3 I = O + D * hit.w;
				// maps (material_shared.h:99-173).  Diffuse map: trilinear fetch at the ray-cone LOD; an alpha texel below 0.5 lets the path through
				float tu = 0, tv = 0;
				if ((mflags & M_ANYMAP) && f.texels) tu = bw * a0.w + bu * a1.w + bv * a2.w, tv = bw * a3.x + bu * a3.y + bv * a3.z;
				else mflags &= ~M_ANYMAP;
				if (mflags & M_DIFFUSEMAP)
				{
					c3 = __ldg( materials + mat * MAT_STRIDE + 3 ), c4 = __ldg( materials + mat * MAT_STRIDE + 4 );
					const float lambda = a3.w + log2f( __ldg( &pass->spreadAngle ) * hit.w * (1.0f / fabsf( dot( D, N ) )) );
					const float4 texel = FetchTexelTrilinear( f.texels, lambda, c3.x * (c3.z + tu), c3.y * (c3.w + tv), __float_as_uint( c2.w ),
						(int)__float_as_uint( c4.x ), (int)__float_as_uint( c4.y ), (int)__float_as_uint( c1.w ) );
					if ((mflags & 2u) && texel.w < 0.5f)
					{
						// pathtracer.h:104-116: the path continues in the same direction from just behind the surface, state unchanged
						if (firstHit && pathLength == maxPathLength) // it ends here: store something sensible, or alpha does not reproject (:143-154)
						{
							const uint32_t pn = PackNormal2( N ) + viaSpecular;
							packFeatures( 0, pn, 0 );
							f.worldPos[pixelIdx] = make_float4( I.x, I.y, I.z, __uint_as_float( pn ) );
						}
						if (pathLength < maxPathLength)
						{
							const V3 po = I + D * f.epsilon;
							eo = po, er = D, en = lastN, et = isfinite( throughput.x + throughput.y + throughput.z ) ? throughput : v3( 0.f, 0.f, 0.f ), epdf = bsdfPdf;
							wantExt = true;
						}
						color = v3( 0.f, 0.f, 0.f ), mflags = 0x80000000u; // nothing else happens at this vertex
					}
					else
					{
						color = color * v3( texel );
						if (mflags & M_DIFFUSEMAP2) // detail layer: additive around 0.5, needs the base map (:120-130)
							color = color + v3( FetchMap( f.texels, materials + mat * MAT_STRIDE + 6, tu, tv ) ) - v3( 0.5f, 0.5f, 0.5f );
					}
				}
				if (mflags & M_NORMALMAP)
				{
					// :132-159; the reference derives the xy scale from bits of the packed record that alias the medium colour (:141),
					// this core takes Vec3Value::scale of CoreMaterial::normals / detailNormals
					const float4* rec = materials + mat * MAT_STRIDE + 8;
					V3 sn = (v3( FetchMap( f.texels, rec, tu, tv ) ) - v3( 0.5f, 0.5f, 0.5f )) * 2.0f;
					const float s0 = __ldg( rec + 1 ).w;
					sn.x *= s0, sn.y *= s0;
					if (mflags & M_NORMALMAP2)
					{
						V3 layer = (v3( FetchMap( f.texels, rec + 2, tu, tv ) ) - v3( 0.5f, 0.5f, 0.5f )) * 2.0f;
						const float s1 = __ldg( rec + 3 ).w;
						layer.x *= s1, layer.y *= s1;
						sn = sn + layer;
					}
					sn = normalize( sn );
					sN = normalize( v3( a4 ) * sn.x + v3( a5 ) * sn.y + sN * sn.z ); // tangent, bitangent as the CoreTri holds them
				}
				if (mflags & M_ROUGHNESSMAP)
				{
					// :161-169: the map's red channel replaces the roughness byte
					roughness = fmaxf( 0.001f, (float)(uint32_t)(FetchMap( f.texels, materials + mat * MAT_STRIDE + 12, tu, tv ).x * 255.0f) * (1.0f / 255.0f) );
					if (DISNEY) dd.roughness = roughness;
				}
				// (FILTERINGCORE, material_shared.h:172-177: no channel of the albedo may be zero, the filter divides by it)
				if (FILTER && !(mflags & 0x80000000u)) color = v3( fmaxf( 0.05f, color.x ), fmaxf( 0.05f, color.y ), fmaxf( 0.05f, color.z ) );
				if (mflags & 0x80000000u) {}
				else
				if (color.x > 1.0f || color.y > 1.0f || color.z > 1.0f)
				{
					if (firstHit)
					{
						// the path ends on a light: finish the feature record with what there is (:193-201)
						const uint32_t pn = PackNormal2( N ) + viaSpecular;
						packFeatures( HDRtoRGB32( color ), pn, mat );
						f.worldPos[pixelIdx] = make_float4( I.x, I.y, I.z, __uint_as_float( pn ) );
						depthDerivatives( I, N );
					}
					// the path ends on a light (pathtracer.h:120-148); lights are single sided
					const float DdotNL = -dot( D, N );
					if (DdotNL > 0)
					{
						V3 c = throughput * color; // (the reference drops the throughput on the specular branch, :129)
						if (pathLength > 1 && !(flags & S_SPECULAR))
						{
							// the previous vertex did NEE: weigh this implicit connection against it (MIS)
							// which light triangle this is: CoreTri::ltriIdx where the mesh came with CoreTris (kept + 1 in the attribute
							// record; pathtracer.h:139 reads it the same way), else a search for (prim, inst) among the area lights
							int ltri = -1;
							float area = 0;
							const int stored = vertexNormals ? __float_as_int( __ldg( vertexNormals + (size_t)(__float_as_uint( info.x ) + prim) * TRI_ATTR_STRIDE + 6 ).y ) - 1 : -1;
							if (stored >= 0 && stored < f.areaCount && __float_as_int( __ldg( f.areaLights + stored * 6 + 3 ).w ) == prim &&
								__float_as_int( __ldg( f.areaLights + stored * 6 + 4 ).w ) == (int)inst) ltri = stored, area = __ldg( f.areaLights + stored * 6 + 1 ).w;
							else for (int i = 0; i < f.areaCount; i++)
								if (__float_as_int( __ldg( f.areaLights + i * 6 + 3 ).w ) == prim && __float_as_int( __ldg( f.areaLights + i * 6 + 4 ).w ) == (int)inst)
									ltri = i, area = __ldg( f.areaLights + i * 6 + 1 ).w;
							const float lightPdf = ltri >= 0 ? (hit.w * hit.w) / (DdotNL * area) : 0.f;    // CalculateLightPDF, lights_shared.h:96-99
							const float pickProb = ltri >= 0 ? LightPickProb( f, ltri, O, lastN, I ) : 0.f;
							const float w = bsdfPdf + lightPdf * pickProb;
							c = w > 0 ? c * (1.0f / w) : v3( 0.f, 0.f, 0.f );
						}
						contribute( ClampAndFix( c, f.clampValue ) );
					}
				}
				else
				{
					if (roughness <= 0.001f || transmission > 0.999f) flags |= S_SPECULAR; else flags &= ~S_SPECULAR; // pathtracer.h:151
					uint32_t seed = WangHash( pathIdx * 17 + __ldg( &pass->R0 ) + (uint32_t)pathLength * 91771u ); // :154; R0 changes per bounce (rendercore.cpp:716)
					if (firstHit)
					{
						if (flags & S_SPECULAR) f.features[pixelIdx].x = HDRtoRGB32( color ); // a mirror: its tint, multiplied into the albedo of the first diffuse vertex behind it (:215-219)
						else
						{
							// the first diffuse vertex: albedo, normal, depth, world position (:221-231).  (The reference negates the shading normal
							// here whichever side was hit -- `flip ? -fN : fN` with flip = +-1 -- and the filter only ever compares these normals
							// with each other; kept.)
							V3 albedo = color;
							if (viaSpecular) albedo = albedo * RGB32toHDR( f.features[pixelIdx].x );
							const uint32_t pn = PackNormal2( sN * -1.0f ) + viaSpecular;
							packFeatures( HDRtoRGB32( albedo ), pn, mat );
							f.worldPos[pixelIdx] = make_float4( I.x, I.y, I.z, __uint_as_float( pn ) );
							depthDerivatives( I, N );
						}
					}
					const float faceDir = dot( D, N ) > 0 ? -1.f : 1.f;
					const V3 gF = N * faceDir, sF = sN * faceDir; // geometric / shading normal on the side the ray came from
					if (faceDir == 1) absorption = v3( 0.f, 0.f, 0.f ); // the ray travelled outside the medium (:158)
					throughput = throughput * (1.0f / bsdfPdf); // postponed pdf (:161)
					// next event estimation (:163-199)
					if (!(flags & S_SPECULAR) && f.areaCount + f.pointCount + f.spotCount + f.dirCount > 0)
					{
						// (the reference draws these two from the blue-noise table for the first TWO samples only, pathtracer.h:168)
						const int sampleIdx = (int)sampleIdxU, bx = (int)(pixelIdx % f.width), by = (int)(pixelIdx / f.width);
						const uint32_t* const blueTable = __ldg( &pass->blue ) ? f.blueTable : nullptr;
						const bool blue = blueTable && sampleIdx < 2;
						const int blueIdx = FILTER ? sampleIdx + (int)__ldg( &pass->blueSlot ) : sampleIdx; // Optix7Filter/kernels/pathtracer.h:248-249
						const float r0 = blue ? BlueNoiseSampler( blueTable, bx, by, blueIdx, 4 + 4 * pathLength ) : RandomFloat( seed );
						const float r1 = blue ? BlueNoiseSampler( blueTable, bx, by, blueIdx, 5 + 4 * pathLength ) : RandomFloat( seed );
						float pickProb, lightPdf;
						V3 lightColor;
						V3 L = RandomPointOnLight( f, r0, r1, I, sF, pickProb, lightPdf, lightColor ) - I;
						const float dist = sqrtf( dot( L, L ) );
						L = L * (1.0f / dist);
						const float NdotL = dot( L, sF );
						if (NdotL > 0 && lightPdf > 0)
						{
							// EvaluateBSDF (lambert.h:58-69): color / pi, pdf = |cos| / pi; scaled by the chance of a diffuse bounce (:186)
							float lbsdfPdf = fabsf( NdotL ) * INVPI;
							V3 sampled = color * (INVPI * roughness);
							if (DISNEY) dd.color = color, dd.transmittance = absorption, sampled = disney::EvaluateBSDF( dd, sN, tangent, D * -1.0f, L, lbsdfPdf ); // unflipped normal (:186)
							if (lbsdfPdf > 0)
							{
								if (DISNEY) se = ClampAndFix( throughput * sampled * lightColor * (NdotL / (pickProb * lightPdf + lbsdfPdf)), f.clampValue );
								else se = ClampAndFix( throughput * color * lightColor * (INVPI * roughness * (NdotL / (pickProb * lightPdf + lbsdfPdf))), f.clampValue );
								SafeOrigin( I.x, I.y, I.z, L.x, L.y, L.z, gF.x, gF.y, gF.z, f.epsilon, so.x, so.y, so.z );
								sl = L, sdist = dist - 2 * f.epsilon;
								wantShadow = true;
							}
						}
					}
					// extension (:201-239)
					const bool enough = (f.bounceCap == 1 && (flags & S_BOUNCED)) || (f.bounceCap == 2 && (flags & S_BOUNCEDTWICE));
					if (firstHit && !enough && pathLength == maxPathLength)
					{
						// the depth cap ends the path before the feature record was finished for good (:289-298)
						const uint32_t pn = PackNormal2( N ) + viaSpecular;
						packFeatures( 0, pn, 0 );
						f.worldPos[pixelIdx] = make_float4( I.x, I.y, I.z, __uint_as_float( pn ) );
					}
					if (!enough && pathLength < maxPathLength)
					{
						const int sampleIdx = (int)sampleIdxU, bx = (int)(pixelIdx % f.width), by = (int)(pixelIdx / f.width);
						const uint32_t* const blueTable = __ldg( &pass->blue ) ? f.blueTable : nullptr;
						const bool blue = blueTable && sampleIdx < 256; // pathtracer.h:212-222
						const int blueIdx = FILTER ? sampleIdx + (int)__ldg( &pass->blueSlot ) : sampleIdx; // :306-307
						const float r3 = blue ? BlueNoiseSampler( blueTable, bx, by, blueIdx, 6 + 4 * pathLength ) : RandomFloat( seed );
						const float r4 = blue ? BlueNoiseSampler( blueTable, bx, by, blueIdx, 7 + 4 * pathLength ) : RandomFloat( seed );
						// SampleBSDF (lambert.h:71-123)
						V3 R = sF, bsdf = v3( 0.f, 0.f, 0.f );
						float newPdf = 1;
						bool specular = true, refracted = false;
						if (DISNEY)
						{
							specular = false;
							dd.color = color, dd.transmittance = absorption;
							bsdf = disney::SampleBSDF( dd, sN, N, tangent, D * -1.0f, hit.w, r3, r4, R, newPdf, specular );
						}
						else if (r4 < transmission)
						{
							// dielectric: Fresnel reflection or refraction, Beer absorption over the distance travelled inside (:84-103)
							const float eio = faceDir < 0 ? 1.0f / eta : eta;
							const float VdotN = -dot( sF, D ); // wo = -D; positive by construction of sF unless the shading normal leans away
							float F = 1.0f;
							{
								float e2 = eio, c = VdotN;
								if (c < 0.0f) e2 = 1.0f / e2, c = fabsf( c );
								const float sinT2 = e2 * e2 * (1.0f - c * c);
								if (sinT2 <= 1.0f)
								{
									const float ldn = sqrtf( 1.0f - sinT2 );
									const float ra = (c - e2 * ldn) / (c + e2 * ldn), rb = (ldn - e2 * c) / (ldn + e2 * c);
									F = 0.5f * (ra * ra + rb * rb);
								}
							}
							const V3 beer = v3( __expf( -absorption.x * hit.w * 2.0f ), __expf( -absorption.y * hit.w * 2.0f ), __expf( -absorption.z * hit.w * 2.0f ) );
							if (r3 < F) R = D - sF * (2 * dot( D, sF ));
							else
							{
								// Refract_L (:45-56): eta = eio, wi = wo = -D
								const float cosI = fabsf( dot( sF, D ) ), sin2I = fmaxf( 0.0f, 1.0f - cosI * cosI ), sin2T = eio * eio * sin2I;
								if (sin2T >= 1) newPdf = 0; // total internal reflection: the reference carries on with a zero throughput, the path ends here
								else R = D * eio + sF * (eio * cosI - sqrtf( 1.0f - sin2T )), refracted = true;
							}
							bsdf = color * beer * (1.0f / fabsf( dot( sF, R ) ));
						}
						else
						{
							const float pReflect = 1 - roughness; // a specular reflection with probability 1 - roughness
							if (r3 < pReflect)
							{
								R = D - sF * (2 * dot( D, sF ));
								bsdf = color * (1.0f / fabsf( dot( sF, R ) ));
							}
							else
							{
								const float r5 = (r3 - pReflect) / (1 - pReflect), r6 = (r4 - transmission) / (1 - transmission);
								const float term1 = TWOPI * r5, term2 = sqrtf( 1 - r6 ); // DiffuseReflectionCosWeighted, tools_shared.h:250-256
								float sn, cs;
								SinCosDet( term1, sn, cs );
								Tangent2World( cs * term2, sn * term2, sqrtf( r6 ), sF.x, sF.y, sF.z, R.x, R.y, R.z );
								R = normalize( R );
								newPdf = fmaxf( 0.0f, dot( R, sF ) ) * INVPI;
								specular = false;
								bsdf = color * INVPI;
							}
						}
						if (!DISNEY && !refracted && dot( gF, R ) <= 0) newPdf = 0; // the sampled direction must leave on the geometric side (:121); refraction returns before this test
						if (newPdf >= 1e-4f) // also false for NaN (:222)
						{
							if (specular) flags |= S_SPECULAR;
							// Russian roulette from the second diffuse vertex on (:225-227)
							const float p = (!f.roulette || (flags & S_SPECULAR) || !(flags & S_BOUNCED)) ? 1.f : fminf( 1.0f, fmaxf( fmaxf( bsdf.x, bsdf.y ), bsdf.z ) );
							if (!(p < RandomFloat( seed )))
							{
								throughput = throughput * (1.0f / p);
								if (!(flags & S_SPECULAR)) flags |= (flags & S_BOUNCED) ? S_BOUNCEDTWICE : S_BOUNCED; else flags |= S_VIASPECULAR;
								SafeOrigin( I.x, I.y, I.z, R.x, R.y, R.z, gF.x, gF.y, gF.z, f.epsilon, eo.x, eo.y, eo.z );
								if (!isfinite( throughput.x + throughput.y + throughput.z )) throughput = v3( 0.f, 0.f, 0.f );
								er = R, en = sF, et = throughput * bsdf * fabsf( dot( sF, R ) ), epdf = newPdf;
								wantExt = true;
							}
						}
					}
				}
			}
			if (pathLength == 1 && !stored) c.slotS[slot] = make_float4( 0.f, 0.f, 0.f, depth );
		}
		uint32_t sslot, eslot;
		blockAppend2<4>( counters + shadowCounter( pathLength ), wantShadow, wantExt, sWarp, sslot, eslot ); // [2 len] shadow, [2 len + 1] = extension of len + 1
		if (wantShadow)
		{
			c.conO[sslot] = make_float4( so.x, so.y, so.z, f.epsilon );
			c.conD[sslot] = make_float4( sl.x, sl.y, sl.z, sdist );
			c.conE[sslot] = make_float4( se.x, se.y, se.z, __uint_as_float( slot ) ); // the connect wave adds to slotC[slot]
		}
		if (wantExt)
		{
			c.rayOout[eslot] = make_float4( eo.x, eo.y, eo.z, f.epsilon );
			c.rayDout[eslot] = make_float4( er.x, er.y, er.z, 1e34f );
			c.stateOut[eslot] = make_float4( et.x, et.y, et.z, epdf );
			c.infoOut[eslot] = make_float4( __uint_as_float( pathIdx | (flags << PATHIDX_BITS) ), en.x, en.y, en.z );
		}
	}
}

// Diagnostic: the Disney BSDF on its own, for the parity test against the reference's disney.h (tests/test_disney_gpu.py).
// in: 28 floats per item -- color, transmittance, three parameter words (bits), eta, iN, N, iT, wow, then (distance, r0, r1) for
// mode 0 = SampleBSDF or wiw for mode 1 = EvaluateBSDF; out: 8 floats -- (wiw, pdf, value, specular) / (-, pdf, value, 0).
__global__ void disneyProbeKernel( const int n, const int mode, const float* __restrict__ in, float* __restrict__ out )
{
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const float* p = in + (size_t)i * 28;
	disney::Data d;
	d.color = v3( p[0], p[1], p[2] ), d.transmittance = v3( p[3], p[4], p[5] );
	disney::Unpack( d, __float_as_uint( p[6] ), __float_as_uint( p[7] ), __float_as_uint( p[8] ) );
	d.eta = p[9];
	disney::Tint( d, d.color );
	const V3 iN = v3( p[10], p[11], p[12] ), N = v3( p[13], p[14], p[15] ), iT = v3( p[16], p[17], p[18] ), wow = v3( p[19], p[20], p[21] );
	V3 wiw = v3( p[22], p[23], p[24] ), value;
	float pdf = 0;
	bool specular = false;
	if (mode == 0) value = disney::SampleBSDF( d, iN, N, iT, wow, p[22], p[23], p[24], wiw, pdf, specular );
	else value = disney::EvaluateBSDF( d, iN, iT, wow, wiw, pdf );
	float* o = out + (size_t)i * 8;
	o[0] = wiw.x, o[1] = wiw.y, o[2] = wiw.z, o[3] = pdf, o[4] = value.x, o[5] = value.y, o[6] = value.z, o[7] = specular ? 1.f : 0.f;
}

// Deterministic accumulation.  The reference adds every contribution straight into accumulator[pixel] (non-atomic RMW, racy for
// spp > 1: connections.h:33); float atomics fix the race but leave the ORDER of a pixel's adds to the scheduler as soon as several
// samples of a pixel are in flight in one pass, so the same frame would differ in the last bits from run to run and from one sharding
// to another.  Here every path owns two slots for the pass -- slotS for what its vertices add (shade waves, in wave order), slotC for
// its next-event contributions (connect waves, in wave order) -- and this kernel folds them into the accumulator, samples in
// ascending order: acc = ((acc + S_0) + C_0) + S_1 ...  The result depends on neither the pass partition, the chain count nor the
// tile sharding: a tile-sharded multi-GPU frame is bit-identical to the single-GPU one at any spp.  With `frame` set the kernel also
// writes the finalized frame (lib/CUDA/shared_kernel_code/finalize_shared.h:29-45) -- the last pass of an unsharded frame.
__global__ void __launch_bounds__( 256 ) accumulateKernel( const FrameParams f, const PassParams* __restrict__ pass, const uint32_t chain, const uint32_t chainTiles,
	const float4* __restrict__ slotS, const float4* __restrict__ slotC, float4* __restrict__ accumulator, float4* __restrict__ frame, const uint32_t indirectOffset )
{
	const uint32_t perSample = chainTiles * TILE_PIXELS, samples = __ldg( &pass->samplesThisPass );
	const float scale = __ldg( &pass->finalScale );
	for (uint32_t q = blockIdx.x * blockDim.x + threadIdx.x; q < perSample; q += gridDim.x * blockDim.x)
	{
		const uint32_t chainTile = q / TILE_PIXELS, inTile = q % TILE_PIXELS;
		const uint32_t tile = (chainTile * f.chainCount + chain) * f.shardCount + f.shardIndex;
		uint32_t tx, ty;
		tileCoords( tile, f.tilesX, f.tileSkew, tx, ty );
		const uint32_t x = tx * TILE_W + (inTile % TILE_W), y = ty * TILE_H + inTile / TILE_W;
		if (x >= f.width || y >= f.height) continue;
		const uint32_t pixel = x + y * f.width;
		for (uint32_t plane = 0; plane < (f.features ? 2u : 1u); plane++) // filter mode: direct light, then indirect (the slots indirectOffset further on)
		{
			const float4 *planeS = slotS + (size_t)plane * indirectOffset, *planeC = slotC + (size_t)plane * indirectOffset;
			float4 a = accumulator[pixel + plane * f.pixelCount];
			for (uint32_t s = 0; s < samples; s++)
			{
				const float4 S = __ldcs( planeS + (size_t)s * perSample + q ), C = __ldcs( planeC + (size_t)s * perSample + q );
				a.x += S.x, a.y += S.y, a.z += S.z, a.w += S.w;
				a.x += C.x, a.y += C.y, a.z += C.z;
			}
			accumulator[pixel + plane * f.pixelCount] = a;
			if (frame && !f.features) frame[pixel] = make_float4( a.x * scale, a.y * scale, a.z * scale, 1.0f );
		}
	}
}

// lib/CUDA/shared_kernel_code/finalize_shared.h:29-45: accumulator * 1/samples -> render target (here: a device buffer)
__global__ void finalizeKernel( const float4* __restrict__ accumulator, float4* __restrict__ frame, const uint32_t pixels, const float scale )
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= pixels) return;
	const float4 a = accumulator[i];
	frame[i] = make_float4( a.x * scale, a.y * scale, a.z * scale, 1.0f );
}


// ---- multi-GPU gather over peer memory (tile-sharded frames) -----------------------------------------------------------------
// Each rank renders its interleaved tiles into ITS OWN accumulator; the other ranks' accumulators are mapped into this process
// (cudaIpcOpenMemHandle: NVLink peer access).  One kernel per frame then pulls every foreign tile straight out of its owner's
// memory, stores it into the local accumulator and writes the finalized frame in the same pass: gather + finalize fused, each
// pixel crosses NVLink once ((G-1)/G of the image per GPU; the all-reduce it replaces moves twice that and needs a separate
// finalize launch).  Ranks meet before and after the pull on flags in peer memory.
constexpr int MAX_PEERS = 16;
struct PeerTable {
	float4* acc[MAX_PEERS];      // acc[r] = rank r's accumulator (own rank: the local pointer)
	uint32_t* flags[MAX_PEERS];  // flags[r] = rank r's arrival words, one per sender
};

// arrival: tell every rank that `rank` has reached `epoch`, then wait until every rank has told us.  One block, one thread per
// rank.  The stores are system-scope releases after a system fence, so what earlier kernels of this stream wrote (the
// accumulator) is visible to a peer that has seen the flag.  A peer that never arrives makes the wait give up after ~2 s and
// raise *timeout, instead of hanging the device (limit: Setting peerTimeoutMs, converted to SM clock ticks by the host); the pull
// kernel that follows does nothing when the flag is up, and the host clears the flag after reporting the error.
// The epoch is a device-resident word of this rank that the kernel itself advances (every rank runs the same sequence of barriers, so the
// words stay in step): no host-side argument changes from frame to frame, which lets the barrier be part of the captured frame graph.
__global__ void peerBarrierKernel( const PeerTable peers, const int world, const int rank, uint32_t* epochWord, uint32_t* timeout, const long long limitTicks )
{
	__shared__ uint32_t sEpoch;
	const int t = threadIdx.x;
	if (t == 0) sEpoch = ++*epochWord;
	__syncthreads();
	const uint32_t epoch = sEpoch;
	if (t >= world) return;
	__threadfence_system();
	asm volatile( "st.release.sys.global.u32 [%0], %1;" ::"l"( peers.flags[t] + rank ), "r"( epoch ) : "memory" );
	const uint32_t* mine = peers.flags[rank] + t;
	const long long t0 = clock64();
	while (true)
	{
		uint32_t v;
		asm volatile( "ld.acquire.sys.global.u32 %0, [%1];" : "=r"( v ) : "l"( mine ) : "memory" );
		if ((int32_t)(v - epoch) >= 0) break;
		if (clock64() - t0 > limitTicks) { atomicExch( timeout, 1u ); break; }
		__nanosleep( 200 );
	}
}

__global__ void gatherFinalizeKernel( const PeerTable peers, const int world, const int rank, float4* __restrict__ frame, const uint32_t width,
	const uint32_t height, const uint32_t tilesX, const uint32_t tileSkew, const float scaleArg, const PassParams* __restrict__ lastPass, const uint32_t* __restrict__ timeout )
{
	if (*timeout) return; // a peer did not arrive: its tiles may be incomplete, nothing is pulled this frame
	const float scale = lastPass ? __ldg( &lastPass->finalScale ) : scaleArg; // inside a frame: 1 / samples, as the frame's last pass has it
	const uint32_t pixels = width * height;
	float4* const own = peers.acc[rank];
	for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < pixels; i += gridDim.x * blockDim.x)
	{
		const uint32_t x = i % width, y = i / width;
		const uint32_t owner = tileIndex( x / TILE_W, y / TILE_H, tilesX, tileSkew ) % (uint32_t)world; // as generateEyeRaysKernel deals the tiles
		float4 a;
		if (owner == (uint32_t)rank) a = own[i];
		else
		{
			// a 32-pixel tile row is 512 contiguous bytes in the owner's memory: full NVLink packets
			const float4* src = peers.acc[owner] + i;
			asm volatile( "ld.global.relaxed.sys.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"( a.x ), "=f"( a.y ), "=f"( a.z ), "=f"( a.w ) : "l"( src ) : "memory" );
			own[i] = a;
		}
		frame[i] = make_float4( a.x * scale, a.y * scale, a.z * scale, 1.0f );
	}
}

} // namespace

constexpr int MAX_CHAINS = 4;
constexpr int COUNTER_STRIDE = (COUNTER_WORDS + 15) & ~15; // one chain's counters of one pass (64-byte aligned)
constexpr int FETCH_BASE = 16;                               // fetch-counter words 16 + 4 chain + {0,1: extend, 2,3: connect}, see enqueueFrame

// One chain of waves (see ChainArgs): buffers for its share of a pass, two streams, the events that order them
struct Chain {
	DeviceBuffer rayO[2], rayD[2], state[2], info[2], hits, conO[2], conD[2], conE[2], slotS, slotC;
	uint32_t capacity = 0; // paths per pass the buffers hold
	cudaStream_t main = nullptr, connect = nullptr;
	cudaEvent_t shadeDone[MAX_PATH_LENGTH + 1] = {}, connectDone[MAX_PATH_LENGTH + 1] = {}, done = nullptr;
	DeviceBuffer* all( int i )
	{
		DeviceBuffer* b[] = { &rayO[0], &rayO[1], &rayD[0], &rayD[1], &state[0], &state[1], &info[0], &info[1], &hits, &conO[0], &conO[1], &conD[0], &conD[1],
			&conE[0], &conE[1], &slotS, &slotC };
		return i < 17 ? b[i] : nullptr;
	}
};

// Everything enqueueFrame reads that is not a stream or an event.  Plain data on purpose: its bytes are the key of the captured frame
// graph -- a frame whose plan equals the previous one's replays the graph, anything else (resolution, spp, lights, a reallocated
// buffer ...) re-enqueues.  What changes every frame without changing the plan (camera, sample numbering) is in PassParams.
// the filter's share of a frame: this frame's reprojection constants and which of the ping-pong buffers plays which part
struct FilterPlan {
	FilterReprojection reprojection;
	float4 *prevWorldPos, *shading, *prevPixels, *taaOut, *moments, *prevMoments, *filteredIn, *filteredOut;
	float2* motion;
	uint32_t taa;
};

struct FramePlan {
	FrameParams f;
	FilterPlan filter;               // used when f.features is set
	TraceArgs trace;                 // scene pools (RenderCore::baseArgs)
	const float4 *shadePool, *vnormPool, *instShadePool, *materials;
	float4 *accumulator, *frame;
	PassParams *passDev, *passHost;
	uint32_t *countersDev, *countersHost, *fetchCounters;
	struct { float4 *rayO[2], *rayD[2], *state[2], *info[2], *hits, *conO[2], *conD[2], *conE[2], *slotS, *slotC; uint32_t tiles, capacity; } chain[MAX_CHAINS];
	uint32_t chains, passes, maxLen, overlap, anyLight, bsdf, clearAccumulator, timedPasses;
	int isect, smCount, shadeGrid;
	float l2Persist;
	// tile-sharded frame with open peers: the gather over peer memory is the tail of the frame (see enqueueFrame)
	uint32_t gather;
	int peerWorld, peerRank;
	PeerTable peers;
	uint32_t *peerEpochWord, *peerTimeout, *peerTimeoutHost;
	long long peerLimitTicks;
};

struct WavefrontState {
	int width = 0, height = 0, spp = 1, maxPathLength = 5;
	int shardIndex = 0, shardCount = 1, sampleOffset = 0, jitter = 1;
	int chains = 0;        // Setting chains: 0 = automatic (by the size of a pass), else 1..MAX_CHAINS
	int useGraph = 1;      // Setting frameGraph: a frame whose plan repeats is replayed as a CUDA graph
	int waveTimers = 1;    // Setting waveTimers: per-wave CUDA events for the CoreStats trace / shade times
	int gatherInRender = 1; // Setting gatherInRender: with open peers, a tile-sharded Render() ends with the peer-memory gather
	uint32_t pathsPerPass = 128u << 20; // (272 B of path buffers each, allocated for what a frame needs: the 4K x 16 spp frame is one pass of 133 M paths and 36 GB; with 32 Mi it ran 1.2 % slower in four)
	Chain chain[MAX_CHAINS];
	DeviceBuffer counters, passParams, accumOwn, frame, materials, lights, skyImage, texels, blueNoise;
	PinnedBuffer hostPass, hostCounters;
	bool haveBlueNoise = false;
	struct TextureRec { uint32_t offset, width, height, mipLevels; bool valid; };
	std::vector<TextureRec> textures;
	float4* accumExternal = nullptr;
	uint32_t materialCount = 0;
	float sky[3] = { 0, 0, 0 }, clampValue = 10.0f; // clampValue: the RenderSystem's default (rendersystem.cpp:228-233 sends it every frame)
	int lightCounts[4] = { 0, 0, 0, 0 };             // area, point, spot, directional; packed in this order in `lights`
	uint32_t skyWidth = 0, skyHeight = 0;
	int bounceCap = 0, roulette = 1, bsdf = 0; // bsdf: 0 = Lambert + speculars, 1 = Disney principled (rendercore_optix7/kernels/bsdf.h picks one at compile time)
	uint32_t samplesTaken = 0, frameCounter = 0;
	int probeX = -1, probeY = -1;
	// Filter mode.  filterCore = 1 makes the core behave as the reference's RenderCore_Optix7Filter: the keys "filter", "TAA", "clampDirect" and
	// "clampIndirect", which the RenderSystem sends to every core every frame (rendersystem.cpp:231-236), are then honoured as that core honours
	// them (rendercore.cpp:648-670; defaults rendercore.h:30-33).  With filterCore = 0, the default, they are accepted and ignored, as in
	// lib/rendercore_optix7.
	int filterCore = 0, filterOn = 1, taaOn = 1;
	float filterClampDirect = 2.5f, filterClampIndirect = 15.0f;
	DeviceBuffer features, worldPos[2], deltaDepth, shading[2], taaOut, motion, moments[2], filtered[2];
	uint32_t filterPixels = 0;       // the size the filter buffers were made (and cleared) for
	int filterFlip = 0;              // which of the pairs is "current" this frame
	uint32_t frameCycle = 0;         // TAA jitter: four sub-pixel offsets in turn (rendercore.cpp:733-737)
	uint32_t blueSlot = 0, filterFrames = 0; // blue-noise slot of the next filtered frame (+ 1 per frame, mod 256); filtered frames rendered so far
	float prevView[12] = {}, prevJ0 = 0, prevJ1 = 0;
	cudaEvent_t filterBegin = nullptr, filterEnd = nullptr;
	cudaEvent_t frameBegin = nullptr, frameEnd = nullptr, fork = nullptr, gatherWait = nullptr, gatherBegin = nullptr, gatherEnd = nullptr;
	std::vector<cudaEvent_t> timers; // per (timed pass, chain, path length): extend begin, extend end, shade end, connect begin, connect end
	int overlapConnect = 1;
	float l2PersistApplied = 0.f;
	const void* l2WindowNodes = nullptr;
	// the captured frame
	cudaStream_t captureStream = nullptr;
	cudaGraphExec_t graphExec = nullptr;
	std::vector<uint8_t> graphKey, lastKey;
	uint32_t graphLaunches = 0, graphCaptures = 0;
	float4* accumulator() { return accumExternal ? accumExternal : (float4*)accumOwn.ptr; }
	// peer-memory gather (OpenPeers): the other ranks' accumulators and arrival flags, mapped through CUDA IPC
	PeerTable peers{};
	int peerWorld = 0, peerRank = 0;
	float peerTimeoutMs = 2000.f;
	void* peerMapped[2 * MAX_PEERS] = {};
	DeviceBuffer peerFlags; // MAX_PEERS arrival words + 1 timeout word, written by the peers
	void closePeers()
	{
		for (void*& p : peerMapped) if (p) { cudaIpcCloseMemHandle( p ); p = nullptr; }
		peerWorld = 0;
	}
	void dropGraph()
	{
		if (graphExec) cudaGraphExecDestroy( graphExec );
		graphExec = nullptr, graphKey.clear(), lastKey.clear();
	}
	// frames on their way to the host (ReadFrameAsync): a device copy of the finished frame, the stream that carries it out, two events
	DeviceBuffer frameStage;
	cudaStream_t readStream = nullptr;
	cudaEvent_t frameCopied = nullptr, frameRead = nullptr;
	bool framePending = false;
	~WavefrontState()
	{
		if (readStream) cudaStreamSynchronize( readStream ), cudaStreamDestroy( readStream );
		if (frameCopied) cudaEventDestroy( frameCopied );
		if (frameRead) cudaEventDestroy( frameRead );
		frameStage.release();
		DeviceBuffer* all[] = { &counters, &passParams, &accumOwn, &frame, &materials, &lights, &skyImage, &texels, &blueNoise, &features, &worldPos[0], &worldPos[1],
			&deltaDepth, &shading[0], &shading[1], &taaOut, &motion, &moments[0], &moments[1], &filtered[0], &filtered[1] };
		closePeers();
		dropGraph();
		peerFlags.release();
		for (DeviceBuffer* b : all) b->release();
		hostPass.release(), hostCounters.release();
		for (Chain& c : chain)
		{
			for (int i = 0; c.all( i ); i++) c.all( i )->release();
			for (cudaEvent_t e : c.shadeDone) if (e) cudaEventDestroy( e );
			for (cudaEvent_t e : c.connectDone) if (e) cudaEventDestroy( e );
			if (c.done) cudaEventDestroy( c.done );
			if (c.main) cudaStreamDestroy( c.main );
			if (c.connect) cudaStreamDestroy( c.connect );
		}
		for (cudaEvent_t e : timers) if (e) cudaEventDestroy( e );
		for (cudaEvent_t e : { frameBegin, frameEnd, fork, gatherWait, gatherBegin, gatherEnd, filterBegin, filterEnd }) if (e) cudaEventDestroy( e );
		if (captureStream) cudaStreamDestroy( captureStream );
	}
};

void DestroyWavefront( WavefrontState* wf ) { delete wf; }

static WavefrontState* state( WavefrontState*& wf )
{
	if (!wf) wf = new WavefrontState();
	return wf;
}

int RenderCore::SettingWavefront( const char* name, float value )
{
	WavefrontState* w = state( wf );
	const int v = (int)value;
	if (!strcmp( name, "maxPathLength" )) { if (v < 1 || v > MAX_PATH_LENGTH) return fail( "Setting maxPathLength: 1..16" ); w->maxPathLength = v; return 0; }
	if (!strcmp( name, "shardCount" )) { if (v < 1) return fail( "Setting shardCount: >= 1" ); w->shardCount = v; if (w->shardIndex >= v) w->shardIndex = 0; return 0; }
	if (!strcmp( name, "shardIndex" )) { if (v < 0 || v >= w->shardCount) return fail( "Setting shardIndex: 0..shardCount-1 (set shardCount first)" ); w->shardIndex = v; return 0; }
	if (!strcmp( name, "sampleOffset" )) { if (v < 0) return fail( "Setting sampleOffset: >= 0" ); w->sampleOffset = v; return 0; }
	if (!strcmp( name, "jitter" )) { w->jitter = value >= 0.5f; return 0; }
	if (!strcmp( name, "clampValue" )) { if (!(value > 0)) return fail( "Setting clampValue: > 0" ); w->clampValue = value; return 0; }
	if (!strcmp( name, "overlapConnect" )) { w->overlapConnect = value >= 0.5f; return 0; }
	if (!strcmp( name, "chains" )) { if (v < 0 || v > MAX_CHAINS) return fail( "Setting chains: 0 (automatic) or 1..4" ); w->chains = v; return 0; }
	if (!strcmp( name, "frameGraph" )) { w->useGraph = value >= 0.5f; if (!w->useGraph) w->dropGraph(); return 0; }
	if (!strcmp( name, "waveTimers" )) { w->waveTimers = value >= 0.5f; return 0; }
	if (!strcmp( name, "gatherInRender" )) { w->gatherInRender = value >= 0.5f; return 0; }
	if (!strcmp( name, "peerTimeoutMs" )) { if (!(value > 0)) return fail( "Setting peerTimeoutMs: > 0" ); w->peerTimeoutMs = value; return 0; }
	if (!strcmp( name, "russianRoulette" )) { w->roulette = value >= 0.5f; return 0; }
	if (!strcmp( name, "filterCore" )) { w->filterCore = value >= 0.5f; return 0; }
	if (!strcmp( name, "filter" )) { w->filterOn = value != 0; return 0; }
	if (!strcmp( name, "TAA" )) { w->taaOn = value != 0; return 0; }
	if (!strcmp( name, "clampDirect" )) { w->filterClampDirect = value; return 0; }
	if (!strcmp( name, "clampIndirect" )) { w->filterClampIndirect = value; return 0; }
	if (!strcmp( name, "bsdf" )) { if (v < 0 || v > 1) return fail( "Setting bsdf: 0 (Lambert, lib/sharedBSDFs/lambert.h) or 1 (Disney, disney.h)" ); w->bsdf = v; return 0; }
	if (!strcmp( name, "diffuseBounceCap" )) { if (v < 0 || v > 2) return fail( "Setting diffuseBounceCap: 0 (off), 1 or 2" ); w->bounceCap = v; return 0; }
	if (!strcmp( name, "pathsPerPass" )) { if (value < 1024) return fail( "Setting pathsPerPass: >= 1024" ); w->pathsPerPass = (uint32_t)value; return 0; }
	return fail( std::string( "Setting: unknown key " ) + name );
}

// CoreAPI_Base::SetTarget (core_api_base.h:90): only width/height/spp are used (lib/rendercore_optix7/rendercore.cpp:286-330)
int RenderCore::SetTarget( int width, int height, int spp )
{
	if (width <= 0 || height <= 0 || spp <= 0) return fail( "SetTarget: width, height and spp must be positive" );
	WavefrontState* w = state( wf );
	const size_t px = (size_t)width * height;
	cudaError_t e;
	// peers have this core's accumulator mapped: if it has to move, the mappings are void on every rank -- close ours; the host
	// layer exchanges handles again (b200_peer_handle / b200_open_peers) before the next b200_gather_finalize
	if (w->peerWorld > 0 && px * 16 > w->accumOwn.bytes) { cudaStreamSynchronize( stream ); w->closePeers(); }
	w->width = width, w->height = height, w->spp = spp;
	if ((e = w->accumOwn.reserve( px * 16, false, stream )) != cudaSuccess || (e = w->frame.reserve( px * 16, false, stream )) != cudaSuccess) return fail( "SetTarget buffers", e );
	if ((e = cudaMemsetAsync( w->accumulator(), 0, px * 16, stream )) != cudaSuccess) return fail( "SetTarget clear", e );
	w->samplesTaken = 0;
	for (cudaEvent_t* ev : { &w->frameBegin, &w->frameEnd, &w->gatherWait, &w->gatherBegin, &w->gatherEnd }) if (!*ev && (e = cudaEventCreate( ev )) != cudaSuccess) return fail( "cudaEventCreate", e );
	if (!w->fork && (e = cudaEventCreateWithFlags( &w->fork, cudaEventDisableTiming )) != cudaSuccess) return fail( "cudaEventCreate", e );
	if (!w->captureStream && (e = cudaStreamCreateWithFlags( &w->captureStream, cudaStreamNonBlocking )) != cudaSuccess) return fail( "capture stream", e );
	for (Chain& c : w->chain)
	{
		for (cudaEvent_t& ev : c.shadeDone) if (!ev && (e = cudaEventCreateWithFlags( &ev, cudaEventDisableTiming )) != cudaSuccess) return fail( "cudaEventCreate", e );
		for (cudaEvent_t& ev : c.connectDone) if (!ev && (e = cudaEventCreateWithFlags( &ev, cudaEventDisableTiming )) != cudaSuccess) return fail( "cudaEventCreate", e );
		if (!c.done && (e = cudaEventCreateWithFlags( &c.done, cudaEventDisableTiming )) != cudaSuccess) return fail( "cudaEventCreate", e );
		if (!c.main && (e = cudaStreamCreateWithFlags( &c.main, cudaStreamNonBlocking )) != cudaSuccess) return fail( "chain stream", e );
		if (!c.connect && (e = cudaStreamCreateWithFlags( &c.connect, cudaStreamNonBlocking )) != cudaSuccess) return fail( "chain stream", e );
	}
	return 0;
}

int RenderCore::SetAccumulatorBuffer( void* devicePtr )
{
	state( wf )->accumExternal = (float4*)devicePtr;
	return 0;
}

// CoreAPI_Base::SetLights (core_api_base.h:102-105): the four arrays in the reference's host layouts (common_classes.h:275-345:
// CoreLightTri 96 B, CorePointLight 32 B, CoreSpotLight 48 B, CoreDirectionalLight 32 B), copied as float4 records
int RenderCore::SetLights( const void* area, int nArea, const void* point, int nPoint, const void* spot, int nSpot, const void* dir, int nDir )
{
	WavefrontState* w = state( wf );
	if (nArea < 0 || nPoint < 0 || nSpot < 0 || nDir < 0 || (nArea && !area) || (nPoint && !point) || (nSpot && !spot) || (nDir && !dir))
		return fail( "SetLights: bad arguments" );
	const size_t bytes[4] = { (size_t)nArea * 96, (size_t)nPoint * 32, (size_t)nSpot * 48, (size_t)nDir * 32 };
	const void* src[4] = { area, point, spot, dir };
	std::vector<char> packed( bytes[0] + bytes[1] + bytes[2] + bytes[3] + 16 );
	size_t off = 0;
	for (int k = 0; k < 4; k++) { if (bytes[k]) memcpy( packed.data() + off, src[k], bytes[k] ); off += bytes[k]; }
	cudaError_t e;
	if ((e = w->lights.reserve( packed.size(), false, stream )) != cudaSuccess) return fail( "lights", e );
	if ((e = cudaMemcpyAsync( w->lights.ptr, packed.data(), packed.size(), cudaMemcpyHostToDevice, stream )) != cudaSuccess) return fail( "lights upload", e );
	if ((e = cudaStreamSynchronize( stream )) != cudaSuccess) return fail( "lights sync", e );
	w->lightCounts[0] = nArea, w->lightCounts[1] = nPoint, w->lightCounts[2] = nSpot, w->lightCounts[3] = nDir;
	return 0;
}

// headless shortcut: exactly one point light (energy = r + g + b as the RenderSystem computes it, host_light.cpp), or none
int RenderCore::SetPointLight( const float* pos, const float* radiance )
{
	if (!pos || !radiance) return SetLights( nullptr, 0, nullptr, 0, nullptr, 0, nullptr, 0 );
	const float rec[8] = { pos[0], pos[1], pos[2], radiance[0] + radiance[1] + radiance[2], radiance[0], radiance[1], radiance[2], 0 };
	return SetLights( nullptr, 0, rec, 1, nullptr, 0, nullptr, 0 );
}

// constant sky
int RenderCore::SetSky( const float* radiance )
{
	if (!radiance) return fail( "SetSky: null" );
	WavefrontState* w = state( wf );
	memcpy( w->sky, radiance, 12 );
	w->skyWidth = w->skyHeight = 0;
	return 0;
}

// CoreAPI_Base::SetSkyData (core_api_base.h:106): float3 pixels[width * height], equirectangular (tools_shared.h:185-192)
int RenderCore::SetSkyData( const float* rgb, int width, int height )
{
	WavefrontState* w = state( wf );
	if (!rgb || width <= 0 || height <= 0) return fail( "SetSkyData: bad arguments" );
	const size_t n = (size_t)width * height;
	std::vector<float> px( n * 4 );
	for (size_t i = 0; i < n; i++) px[i * 4] = rgb[i * 3], px[i * 4 + 1] = rgb[i * 3 + 1], px[i * 4 + 2] = rgb[i * 3 + 2], px[i * 4 + 3] = 1.0f;
	cudaError_t e;
	if ((e = w->skyImage.reserve( n * 16, false, stream )) != cudaSuccess) return fail( "sky image", e );
	if ((e = cudaMemcpyAsync( w->skyImage.ptr, px.data(), n * 16, cudaMemcpyHostToDevice, stream )) != cudaSuccess) return fail( "sky upload", e );
	if ((e = cudaStreamSynchronize( stream )) != cudaSuccess) return fail( "sky sync", e );
	w->skyWidth = (uint32_t)width, w->skyHeight = (uint32_t)height;
	return 0;
}

// CoreAPI_Base::SetTextures (core_api_base.h:98): CoreTexDesc[count] in the host layout (common_classes.h:246-269, 40 bytes: data
// pointer, width, height, flags, pixelCount, firstPixel, MIPlevels, storage).  ARGB32 textures are copied, MIP levels included, into
// one texel pool (what SyncStorageType does, lib/rendercore_optix7/rendercore.cpp:447-505), NRM32 normal maps beside them (the
// reference keeps a second pool with the same texel format) and ARGB128 (HDR) textures as float4 texels at 16-byte aligned places of the
// same pool (see LoadTexel).
int RenderCore::SetTextures( const void* coreTexDescs, int count )
{
	WavefrontState* w = state( wf );
	if (count < 0 || (count > 0 && !coreTexDescs)) return fail( "SetTextures: bad arguments" );
	w->textures.assign( (size_t)count, WavefrontState::TextureRec{ 0, 0, 0, 0, false } );
	stats.argb32TexelCount = stats.argb128TexelCount = stats.nrm32TexelCount = 0;
	std::vector<uint32_t> pool;
	for (int i = 0; i < count; i++)
	{
		const char* d = (const char*)coreTexDescs + (size_t)i * 40;
		const uint32_t* data;
		uint32_t width, height, pixelCount, mips, storage;
		memcpy( &data, d, 8 ), memcpy( &width, d + 8, 4 ), memcpy( &height, d + 12, 4 ), memcpy( &pixelCount, d + 20, 4 ), memcpy( &mips, d + 28, 4 ), memcpy( &storage, d + 32, 4 );
		if (storage > 2 || !data || !width || !height) continue;
		const bool hdr = storage == 1; /* ARGB128: float4 texels; ARGB32 (0) and NRM32 (2): four bytes per texel */
		// the levels actually present: level k has (w >> k)(h >> k) texels
		uint32_t have = 0, need = 0, ww = width, hh = height;
		while (have < (mips ? mips : 1) && ww && hh && need + ww * hh <= pixelCount) need += ww * hh, ww >>= 1, hh >>= 1, have++;
		if (!have) return fail( "SetTextures: pixelCount smaller than width * height" );
		if (hdr)
		{
			pool.resize( (pool.size() + 3) & ~(size_t)3, 0u );
			if (pool.size() / 4 + need >= TEXEL_HDR) return fail( "SetTextures: texel pool too large" );
			w->textures[i] = { (uint32_t)(pool.size() / 4) | TEXEL_HDR, width, height, have, true };
			pool.insert( pool.end(), data, data + (size_t)need * 4 );
			stats.argb128TexelCount += need;
		}
		else
		{
			w->textures[i] = { (uint32_t)pool.size(), width, height, have, true };
			pool.insert( pool.end(), data, data + need );
			(storage == 0 ? stats.argb32TexelCount : stats.nrm32TexelCount) += need;
		}
	}
	if (pool.size() >= TEXEL_HDR) return fail( "SetTextures: texel pool too large" );
	cudaError_t e;
	if ((e = w->texels.reserve( pool.size() * 4 + 64, false, stream )) != cudaSuccess) return fail( "texel pool", e );
	if (!pool.empty() && (e = cudaMemcpyAsync( w->texels.ptr, pool.data(), pool.size() * 4, cudaMemcpyHostToDevice, stream )) != cudaSuccess) return fail( "texel upload", e );
	if ((e = cudaStreamSynchronize( stream )) != cudaSuccess) return fail( "texel sync", e );
	return 0;
}

// CoreAPI_Base::SetMaterials (core_api_base.h:100; after SetTextures, :99): CoreMaterial[count], 688 bytes each
// (common_classes.h:177-238), read through the restated struct of include/lh2_abi.h (layout pinned there by static_asserts:
// float2 is 8-byte aligned in the reference, common_types.h:62, so a Vec3Value keeps uvscale at +24 and uvoffset at +32).
// The Lambert stage keeps color (value + map), flags (1 = SMOOTH, 2 = HASALPHA, host_material.h:66-71), absorption.value (stored as
// 1 - absorption like rendercore_optix7/rendercore.cpp:529-531), roughness, transmission and eta; the other maps are detailColor,
// normals, detailNormals and the roughness map.
int RenderCore::SetMaterials( const void* coreMaterials, int count )
{
	if (count < 0 || (count > 0 && !coreMaterials)) return fail( "SetMaterials: bad arguments" );
	std::vector<float> table( (size_t)count * 16, 0.f );
	std::vector<uint32_t> words( (size_t)count * 3, 0u );
	std::vector<MaterialMap> maps( (size_t)count * 4 );
	for (int i = 0; i < count; i++)
	{
		lh2abi::CoreMaterial m;
		memcpy( &m, (const char*)coreMaterials + (size_t)i * sizeof( m ), sizeof( m ) ); // the caller's array need not be aligned for us
		const lh2abi::CoreMaterial::Vec3Value* vmaps[3] = { &m.detailColor, &m.normals, &m.detailNormals };
		for (int k = 0; k < 3; k++)
		{
			MaterialMap& mm = maps[(size_t)i * 4 + k];
			mm.texture = vmaps[k]->textureID, mm.scale = vmaps[k]->scale;
			mm.uvscale[0] = vmaps[k]->uvscale.x, mm.uvscale[1] = vmaps[k]->uvscale.y, mm.uvoffset[0] = vmaps[k]->uvoffset.x, mm.uvoffset[1] = vmaps[k]->uvoffset.y;
		}
		{
			MaterialMap& mm = maps[(size_t)i * 4 + 3];
			mm.texture = m.roughness.textureID, mm.scale = m.roughness.scale;
			mm.uvscale[0] = m.roughness.uvscale.x, mm.uvscale[1] = m.roughness.uvscale.y, mm.uvoffset[0] = m.roughness.uvoffset.x, mm.uvoffset[1] = m.roughness.uvoffset.y;
		}
		float* t = &table[(size_t)i * 16];
		t[0] = m.color.value.x, t[1] = m.color.value.y, t[2] = m.color.value.z, t[3] = m.roughness.value;
		t[4] = m.transmission.value, t[5] = m.eta.value, t[6] = (float)(m.flags & 3u);
		t[8] = 1 - m.absorption.value.x, t[9] = 1 - m.absorption.value.y, t[10] = 1 - m.absorption.value.z;
		t[11] = (float)m.color.textureID;
		t[12] = m.color.uvscale.x, t[13] = m.color.uvscale.y, t[14] = m.color.uvoffset.x, t[15] = m.color.uvoffset.y;
		// the Disney parameters as 0.8 fixed point, four to a word (TOUINT4, rendercore_optix7/rendercore.cpp:511-512, 532-534):
		// metallic, subsurface, specular, roughness | specularTint, anisotropic, sheen, sheenTint | clearcoat, clearcoatGloss, transmission
		const float disney[11] = { m.metallic.value, m.subsurface.value, m.specular.value, m.roughness.value, m.specularTint.value, m.anisotropic.value,
			m.sheen.value, m.sheenTint.value, m.clearcoat.value, m.clearcoatGloss.value, m.transmission.value };
		for (int k = 0; k < 11; k++) words[(size_t)i * 3 + k / 4] += (uint32_t)(disney[k] * 255.0f) << (8 * (k & 3));
	}
	return uploadMaterials( table.data(), words.data(), maps.data(), count );
}

// headless shortcut: (albedo rgb, roughness) per material, opaque, flat normals, no map
int RenderCore::SetMaterialColors( const float* rgba, int count )
{
	if (count < 0 || (count > 0 && !rgba)) return fail( "SetMaterialColors: bad arguments" );
	std::vector<float> table( (size_t)count * 16, 0.f );
	for (int i = 0; i < count; i++)
	{
		float* t = &table[(size_t)i * 16];
		memcpy( t, rgba + (size_t)i * 4, 16 );
		t[5] = 1.f, t[11] = -1.f, t[12] = t[13] = 1.f;
	}
	return SetMaterialTable( table.data(), count );
}

// 16 floats per material: albedo rgb, roughness | transmission, eta, flags (1 SMOOTH + 2 HASALPHA), 0 | extinction rgb, diffuse
// texture index (-1 = none) | uv scale xy, uv offset xy.  Texture indices refer to the last SetTextures.  The Disney parameters
// this table does not carry take the HostMaterial defaults (specular 0.5, the rest 0); roughness and transmission are quantised.
int RenderCore::SetMaterialTable( const float* props16, int count )
{
	if (count < 0 || (count > 0 && !props16)) return fail( "SetMaterialTable: bad arguments" );
	std::vector<uint32_t> words( (size_t)count * 3, 0u );
	for (int i = 0; i < count; i++)
	{
		const float* t = props16 + (size_t)i * 16;
		words[(size_t)i * 3] = ((uint32_t)(0.5f * 255.0f) << 16) + ((uint32_t)(fminf( fmaxf( t[3], 0.f ), 1.f ) * 255.0f) << 24);
		words[(size_t)i * 3 + 2] = (uint32_t)(fminf( fmaxf( t[4], 0.f ), 1.f ) * 255.0f) << 16;
	}
	return uploadMaterials( props16, words.data(), nullptr, count );
}

int RenderCore::uploadMaterials( const float* props16, const uint32_t* disneyWords, const MaterialMap* maps, int count )
{
	WavefrontState* w = state( wf );
	constexpr int REC = MAT_STRIDE * 4; // floats per device record
	std::vector<float> dev( (size_t)count * REC, 0.f );
	for (int i = 0; i < count; i++)
	{
		const float* t = props16 + (size_t)i * 16;
		float* d = &dev[(size_t)i * REC];
		memcpy( d, t, 64 );
		uint32_t flags = (uint32_t)t[6] & 3u, offset = 0, dims[4] = { 0, 0, 0, 0 }, mips = 0;
		const int tex = (int)t[11];
		if (tex >= 0 && tex < (int)w->textures.size() && w->textures[tex].valid)
			flags |= M_DIFFUSEMAP, offset = w->textures[tex].offset, dims[0] = w->textures[tex].width, dims[1] = w->textures[tex].height, mips = w->textures[tex].mipLevels;
		// detail colour, normal map, detail normal map, roughness map: two float4 each from float 24 on; the detail layers need their base map
		static const uint32_t mapBit[4] = { M_DIFFUSEMAP2, M_NORMALMAP, M_NORMALMAP2, M_ROUGHNESSMAP };
		for (int k = 0; maps && k < 4; k++)
		{
			const MaterialMap& mm = maps[(size_t)i * 4 + k];
			if (mm.texture < 0 || mm.texture >= (int)w->textures.size() || !w->textures[mm.texture].valid) continue;
			if ((k == 0 && !(flags & M_DIFFUSEMAP)) || (k == 2 && !(flags & M_NORMALMAP))) continue;
			flags |= mapBit[k];
			float* r = d + 24 + 8 * k;
			const WavefrontState::TextureRec& t = w->textures[mm.texture];
			r[0] = mm.uvscale[0], r[1] = mm.uvscale[1], r[2] = mm.uvoffset[0], r[3] = mm.uvoffset[1];
			memcpy( r + 4, &t.offset, 4 ), memcpy( r + 5, &t.width, 4 ), memcpy( r + 6, &t.height, 4 ), r[7] = mm.scale;
		}
		memcpy( d + 6, &flags, 4 ), memcpy( d + 7, &mips, 4 ), memcpy( d + 11, &offset, 4 ), memcpy( d + 16, dims, 16 );
		memcpy( d + 20, disneyWords + (size_t)i * 3, 12 );
	}
	cudaError_t e;
	if ((e = w->materials.reserve( (size_t)(count + 1) * REC * 4, false, stream )) != cudaSuccess) return fail( "materials", e );
	if (count && (e = cudaMemcpyAsync( w->materials.ptr, dev.data(), (size_t)count * REC * 4, cudaMemcpyHostToDevice, stream )) != cudaSuccess) return fail( "materials upload", e );
	if ((e = cudaStreamSynchronize( stream )) != cudaSuccess) return fail( "materials sync", e );
	w->materialCount = (uint32_t)count;
	return 0;
}

// The blue-noise table of the reference's cores (rendercore_optix7/rendercore.cpp:258-267: 65536 Sobol entries, then the scrambling
// and ranking tiles at word offsets 65536 and 3 * 65536; 5 * 65536 words).  The reference compiles the data in (core_settings.h:207);
// this core takes it from the host.  words = null switches back to xor32 numbers.
int RenderCore::SetBlueNoise( const uint32_t* words, int count )
{
	WavefrontState* w = state( wf );
	if (!words) { w->haveBlueNoise = false; return 0; }
	if (count != 5 * 65536) return fail( "SetBlueNoise: the table has 5 * 65536 words" );
	cudaError_t e;
	if ((e = w->blueNoise.reserve( (size_t)count * 4, false, stream )) != cudaSuccess) return fail( "blue noise table", e );
	if ((e = cudaMemcpyAsync( w->blueNoise.ptr, words, (size_t)count * 4, cudaMemcpyHostToDevice, stream )) != cudaSuccess) return fail( "blue noise upload", e );
	if ((e = cudaStreamSynchronize( stream )) != cudaSuccess) return fail( "blue noise sync", e );
	w->haveBlueNoise = true;
	return 0;
}

int RenderCore::DisneyProbe( int mode, int n, const float* in28, float* out8 )
{
	if (n < 0 || (mode != 0 && mode != 1) || (n > 0 && (!in28 || !out8))) return fail( "DisneyProbe: bad arguments" );
	if (n == 0) return 0;
	DeviceBuffer in, out;
	cudaError_t e;
	if ((e = in.reserve( (size_t)n * 112, false, stream )) != cudaSuccess || (e = out.reserve( (size_t)n * 32, false, stream )) != cudaSuccess) return fail( "DisneyProbe buffers", e );
	e = cudaMemcpyAsync( in.ptr, in28, (size_t)n * 112, cudaMemcpyHostToDevice, stream );
	if (e == cudaSuccess) disneyProbeKernel<<<(n + 127) / 128, 128, 0, stream>>>( n, mode, (const float*)in.ptr, (float*)out.ptr ), e = cudaGetLastError();
	if (e == cudaSuccess) e = cudaMemcpyAsync( out8, out.ptr, (size_t)n * 32, cudaMemcpyDeviceToHost, stream );
	if (e == cudaSuccess) e = cudaStreamSynchronize( stream );
	in.release(), out.release();
	return e == cudaSuccess ? 0 : fail( "DisneyProbe", e );
}

int RenderCore::SetProbe( int x, int y )
{
	state( wf )->probeX = x, state( wf )->probeY = y;
	return 0;
}

typedef void ( *ShadeKernel )( const FrameParams, const PassParams*, const int, const int, const ChainArgs, const float4*, const float4*, const float4*, const float4* );
static ShadeKernel pickShade( const bool disney, const bool filter )
{
	if (filter) return disney ? shadeKernel<true, true> : shadeKernel<false, true>;
	return disney ? shadeKernel<true, false> : shadeKernel<false, false>;
}

// ---- one frame, enqueued ------------------------------------------------------------------------------------------------
// Replaces the host-driven wave loop of the reference (lib/rendercore_optix7/rendercore.cpp:692-748: per bounce a launch, a counter
// read-back and a synchronisation) by ONE enqueue of the whole frame -- all passes, all chains -- with no host involvement between
// its first and its last operation:
//   origin:  upload PassParams, clear counters (+ accumulator on Restart) --fork--> chains --join--> counter download
//   chain c: per pass  generate -> [extend(len) -> shade(len) -> (connect(len) on the chain's second stream)]* -> accumulate
// Wave sizes live in device counters (TraceArgs::countPtr), grids are persistent, so nothing in the sequence depends on a value the
// host would have to read; that also makes the sequence capturable as a CUDA graph (`capturing`: timer events become event-record
// nodes).  Returns the number of kernels enqueued through *launches.
static cudaError_t enqueueFrame( WavefrontState& w, const FramePlan& p, cudaStream_t origin, const bool capturing, uint32_t* launches )
{
	cudaError_t e;
	uint32_t n = 0;
	const unsigned recFlags = capturing ? cudaEventRecordExternal : cudaEventRecordDefault;
	auto timer = [&]( uint32_t pass, uint32_t c, uint32_t len, int k, cudaStream_t s ) {
		if (pass < p.timedPasses) cudaEventRecordWithFlags( w.timers[((size_t)(pass * p.chains + c) * p.maxLen + (len - 1)) * 5 + k], s, recFlags );
	};
	if ((e = cudaMemcpyAsync( p.passDev, p.passHost, sizeof( PassParams ) * p.passes, cudaMemcpyHostToDevice, origin )) != cudaSuccess) return e;
	if ((e = cudaMemsetAsync( p.countersDev, 0, sizeof( uint32_t ) * COUNTER_STRIDE * p.chains * p.passes, origin )) != cudaSuccess) return e;
	if ((e = cudaMemsetAsync( p.fetchCounters + FETCH_BASE, 0, sizeof( uint32_t ) * 4 * MAX_CHAINS, origin )) != cudaSuccess) return e;
	if (p.clearAccumulator && (e = cudaMemsetAsync( p.accumulator, 0, (size_t)p.f.pixelCount * 16 * (p.f.features ? 2 : 1), origin )) != cudaSuccess) return e;
	cudaEventRecordWithFlags( w.frameBegin, origin, recFlags );
	cudaEventRecord( w.fork, origin );
	for (uint32_t c = 0; c < p.chains; c++) cudaStreamWaitEvent( w.chain[c].main, w.fork, 0 );
	const bool overlap = p.overlap && p.anyLight;
	for (uint32_t pass = 0; pass < p.passes; pass++) for (uint32_t c = 0; c < p.chains; c++)
	{
		Chain& ch = w.chain[c];
		const auto& pc = p.chain[c];
		if (!pc.tiles) continue;
		cudaStream_t s = ch.main;
		const PassParams* passDev = p.passDev + pass;
		uint32_t* counters = p.countersDev + (size_t)(pass * p.chains + c) * COUNTER_STRIDE;
		uint32_t* fetch = p.fetchCounters + FETCH_BASE + 4 * c;
		ChainArgs ca;
		memset( &ca, 0, sizeof( ca ) );
		ca.counters = counters, ca.chain = c, ca.chainTiles = pc.tiles, ca.slotS = pc.slotS, ca.slotC = pc.slotC;
		ca.indirectOffset = p.f.features ? pc.capacity : 0;
		ca.rayOout = pc.rayO[0], ca.rayDout = pc.rayD[0], ca.stateOut = pc.state[0], ca.infoOut = pc.info[0];
		generateEyeRaysKernel<<<p.smCount * 4, 256, 0, s>>>( p.f, passDev, ca );
		n++;
		int in = 0;
		for (uint32_t len = 1; len <= p.maxLen; len++)
		{
			const int out = in ^ 1, con = (int)(len & 1);
			// the fetch counter of a wave kernel is zeroed by the wave before it on the same stream (TraceArgs::fetchCounterNext):
			// two words per stream, used alternately, all zero when the frame starts
			const uint32_t par = (pass * p.maxLen + len) & 1;
			TraceArgs a = p.trace;
			a.rayO = pc.rayO[in], a.rayD = pc.rayD[in], a.hits = pc.hits, a.count = pc.capacity, a.countPtr = counters + extCounter( (int)len );
			a.fetchCounter = fetch + par, a.fetchCounterNext = fetch + (par ^ 1);
			timer( pass, c, len, 0, s );
			if ((e = LaunchExtend( a, (IsectMode)p.isect, false, p.smCount, s )) != cudaSuccess) return e;
			timer( pass, c, len, 1, s );
			// shade( len ) writes the shadow rays of set len & 1: connect( len - 2 ) must have read them
			if (overlap && len >= 3) cudaStreamWaitEvent( s, ch.connectDone[len - 2], 0 );
			ca.rayO = pc.rayO[in], ca.rayD = pc.rayD[in], ca.state = pc.state[in], ca.info = pc.info[in], ca.hits = pc.hits;
			ca.rayOout = pc.rayO[out], ca.rayDout = pc.rayD[out], ca.stateOut = pc.state[out], ca.infoOut = pc.info[out];
			ca.conO = pc.conO[con], ca.conD = pc.conD[con], ca.conE = pc.conE[con];
			pickShade( p.bsdf != 0, p.f.features != nullptr )<<<p.shadeGrid, 128, 0, s>>>( p.f, passDev, (int)len, (int)p.maxLen, ca, p.shadePool, p.vnormPool,
				p.instShadePool, p.materials );
			timer( pass, c, len, 2, s );
			n += 2;
			if (p.anyLight)
			{
				// connect( len ): shadow wave fused with the add of the unoccluded potentials into the paths' slotC.  On the chain's second
				// stream it runs beside extend( len + 1 ); nothing waits for it before the end of the pass
				cudaStream_t cs = overlap ? ch.connect : s;
				if (overlap) { cudaEventRecord( ch.shadeDone[len], s ); cudaStreamWaitEvent( cs, ch.shadeDone[len], 0 ); }
				TraceArgs b = p.trace;
				b.rayO = pc.conO[con], b.rayD = pc.conD[con], b.potentials = pc.conE[con], b.accumulator = pc.slotC, b.count = pc.capacity;
				b.countPtr = counters + shadowCounter( (int)len );
				b.fetchCounter = fetch + 2 + par, b.fetchCounterNext = fetch + 2 + (par ^ 1);
				timer( pass, c, len, 3, cs );
				if ((e = LaunchShadow( b, (IsectMode)p.isect, p.smCount, cs )) != cudaSuccess) return e;
				timer( pass, c, len, 4, cs );
				if (overlap) cudaEventRecord( ch.connectDone[len], cs );
				n++;
			}
			in = out;
		}
		if (overlap) cudaStreamWaitEvent( s, ch.connectDone[p.maxLen], 0 ); // the connect stream is in order: the last wave's event covers them all
		const uint32_t perSample = pc.tiles * TILE_PIXELS;
		const uint32_t blocks = std::min( (perSample + 255u) / 256u, (uint32_t)p.smCount * 8u );
		accumulateKernel<<<blocks, 256, 0, s>>>( p.f, passDev, c, pc.tiles, pc.slotS, pc.slotC, p.accumulator, pass + 1 == p.passes ? p.frame : nullptr, ca.indirectOffset );
		n++;
	}
	for (uint32_t c = 0; c < p.chains; c++) { cudaEventRecord( w.chain[c].done, w.chain[c].main ); cudaStreamWaitEvent( origin, w.chain[c].done, 0 ); }
	if (p.f.features)
	{
		// the reconstruction filter and TAA (lib/RenderCore_Optix7Filter/rendercore.cpp:827-861) on the frame's two accumulator planes
		const FilterPlan& fp = p.filter;
		FilterFrame ff;
		ff.accumulator = p.accumulator, ff.features = p.f.features, ff.worldPos = p.f.worldPos, ff.prevWorldPos = fp.prevWorldPos, ff.deltaDepth = p.f.deltaDepth;
		ff.shading = fp.shading, ff.motion = fp.motion, ff.moments = fp.moments, ff.prevMoments = fp.prevMoments, ff.width = (int)p.f.width, ff.height = (int)p.f.height;
		cudaEventRecordWithFlags( w.filterBegin, origin, recFlags );
		if ((e = LaunchFilterPrepare( ff, fp.reprojection, origin )) != cudaSuccess) return e;
		if ((e = LaunchFilterReconstruct( ff, fp.shading, fp.filteredIn, fp.filteredOut, 1, 0, origin )) != cudaSuccess) return e; // filteredIn: the previous frame's first-pass output
		if ((e = LaunchFilterReconstruct( ff, fp.filteredOut, nullptr, fp.filteredIn, 2, 0, origin )) != cudaSuccess) return e;
		if ((e = LaunchFilterReconstruct( ff, fp.filteredIn, nullptr, fp.shading, 3, 1, origin )) != cudaSuccess) return e;
		if (fp.taa)
		{
			if ((e = LaunchFilterTaa( fp.shading, fp.prevPixels, fp.motion, fp.taaOut, ff.width, ff.height, origin )) != cudaSuccess) return e;
			if ((e = LaunchFilterUnsharpen( fp.taaOut, p.frame, ff.width, ff.height, origin )) != cudaSuccess) return e;
		}
		else if ((e = LaunchFilterFinishNoTaa( fp.shading, p.frame, ff.width, ff.height, origin )) != cudaSuccess) return e;
		cudaEventRecordWithFlags( w.filterEnd, origin, recFlags );
		n += fp.taa ? 6 : 5;
	}
	if (p.gather)
	{
		// the exchange step of a tile-sharded frame, as the tail of the frame: arrive (everyone's tiles are accumulated), pull the foreign
		// tiles over NVLink + finalize, arrive again (everyone has pulled: the accumulators may be written by the next frame)
		cudaEventRecordWithFlags( w.gatherWait, origin, recFlags );
		peerBarrierKernel<<<1, 32, 0, origin>>>( p.peers, p.peerWorld, p.peerRank, p.peerEpochWord, p.peerTimeout, p.peerLimitTicks );
		cudaEventRecordWithFlags( w.gatherBegin, origin, recFlags );
		gatherFinalizeKernel<<<p.smCount * 8, 256, 0, origin>>>( p.peers, p.peerWorld, p.peerRank, p.frame, p.f.width, p.f.height, p.f.tilesX, p.f.tileSkew, 0.f,
			p.passDev + (p.passes - 1), p.peerTimeout );
		cudaEventRecordWithFlags( w.gatherEnd, origin, recFlags );
		peerBarrierKernel<<<1, 32, 0, origin>>>( p.peers, p.peerWorld, p.peerRank, p.peerEpochWord, p.peerTimeout, p.peerLimitTicks );
		if ((e = cudaMemcpyAsync( p.peerTimeoutHost, p.peerTimeout, 4, cudaMemcpyDeviceToHost, origin )) != cudaSuccess) return e;
		n += 3;
	}
	if ((e = cudaMemcpyAsync( p.countersHost, p.countersDev, sizeof( uint32_t ) * COUNTER_STRIDE * p.chains * p.passes, cudaMemcpyDeviceToHost, origin )) != cudaSuccess) return e;
	cudaEventRecordWithFlags( w.frameEnd, origin, recFlags );
	if (launches) *launches = n;
	return cudaGetLastError();
}

// CoreAPI_Base::Render (core_api_base.h:94).  converge: 0 = Converge (keep accumulating), 1 = Restart (clear first)
// (common_classes.h:38-42; lib/rendercore_optix7/rendercore.cpp:627-634).
int RenderCore::Render( const void* viewPyramid, int converge )
{
	WavefrontState* w = wf;
	if (!w || w->width == 0) return fail( "Render: SetTarget has not been called" );
	if (!viewPyramid) return fail( "Render: null view" );
	const auto hostT0 = std::chrono::steady_clock::now();
	auto hostMs = [&]() { return std::chrono::duration<float, std::milli>( std::chrono::steady_clock::now() - hostT0 ).count(); };
	if (ensureToplevel()) return 1;
	cudaError_t e;
	const float* vp = (const float*)viewPyramid; // pos, p1, p2, p3 (4 x float3), then 5 floats (common_classes.h:362-385)
	FramePlan plan;
	memset( &plan, 0, sizeof( plan ) ); // (padding included: the plan's bytes are compared)
	FrameParams& f = plan.f;
	memcpy( f.sky, w->sky, 12 );
	f.texels = w->texels.bytes ? (const uint32_t*)w->texels.ptr : nullptr;
	f.clampValue = w->clampValue, f.bounceCap = (uint32_t)w->bounceCap, f.roulette = (uint32_t)w->roulette;
	f.areaCount = w->lightCounts[0], f.pointCount = w->lightCounts[1], f.spotCount = w->lightCounts[2], f.dirCount = w->lightCounts[3];
	f.areaLights = (const float4*)w->lights.ptr, f.pointLights = f.areaLights + f.areaCount * 6, f.spotLights = f.pointLights + f.pointCount * 2;
	f.dirLights = f.spotLights + f.spotCount * 3;
	plan.anyLight = f.areaCount + f.pointCount + f.spotCount + f.dirCount > 0;
	plan.overlap = (uint32_t)w->overlapConnect;
	if (w->skyWidth) f.skyPixels = (const float4*)w->skyImage.ptr, f.skyWidth = w->skyWidth, f.skyHeight = w->skyHeight;
	f.blueTable = w->haveBlueNoise ? (const uint32_t*)w->blueNoise.ptr : nullptr;
	f.epsilon = epsilon, f.width = (uint32_t)w->width, f.height = (uint32_t)w->height, f.pixelCount = f.width * f.height;
	f.tilesX = (f.width + TILE_W - 1) / TILE_W, f.tilesY = (f.height + TILE_H - 1) / TILE_H;
	const uint32_t nTiles = f.tilesX * f.tilesY;
	f.shardIndex = (uint32_t)w->shardIndex, f.shardCount = (uint32_t)w->shardCount, f.tileSkew = tileSkewFor( f.shardCount );
	const uint32_t ownTiles = nTiles > f.shardIndex ? (nTiles - f.shardIndex + f.shardCount - 1) / f.shardCount : 0;
	f.jitter = (uint32_t)w->jitter, f.materialCount = w->materialCount;
	// ---- filter mode: the frame as lib/RenderCore_Optix7Filter renders it (rendercore.cpp:677-742): accumulator cleared every frame, paths stop
	//      after one diffuse bounce, no Russian roulette, pinhole camera through the pixel corners of a view shifted by the TAA jitter
	const bool filtering = w->filterCore && w->filterOn;
	float j0 = 0, j1 = 0;
	if (filtering)
	{
		if (f.shardCount != 1 || w->accumExternal) return fail( "Render: the filter needs the whole frame on one core (shardCount 1, own accumulator)" );
		if (w->peerWorld > 0) return fail( "Render: close the peers before switching the filter on (the accumulator gets a second plane)" );
		const size_t px = f.pixelCount;
		if ((e = w->accumOwn.reserve( px * 32, false, stream )) != cudaSuccess) return fail( "accumulator planes", e );
		DeviceBuffer* sixteen[] = { &w->features, &w->worldPos[0], &w->worldPos[1], &w->deltaDepth, &w->shading[0], &w->shading[1], &w->taaOut, &w->moments[0], &w->moments[1],
			&w->filtered[0], &w->filtered[1] };
		for (DeviceBuffer* b : sixteen) if ((e = b->reserve( px * 16, false, stream )) != cudaSuccess) return fail( "filter buffers", e );
		if ((e = w->motion.reserve( px * 8, false, stream )) != cudaSuccess) return fail( "filter buffers", e );
		if (w->filterPixels != f.pixelCount)
		{
			// a new size: no history (history length 0 in features.w, previous positions nowhere near)
			for (DeviceBuffer* b : sixteen) if ((e = cudaMemsetAsync( b->ptr, 0, px * 16, stream )) != cudaSuccess) return fail( "filter buffers", e );
			cudaMemsetAsync( w->motion.ptr, 0, px * 8, stream );
			w->filterPixels = f.pixelCount, w->frameCycle = 0, w->prevJ0 = w->prevJ1 = 0;
			memcpy( w->prevView, vp, 48 );
		}
		if (!w->filterBegin && ((e = cudaEventCreate( &w->filterBegin )) != cudaSuccess || (e = cudaEventCreate( &w->filterEnd )) != cudaSuccess)) return fail( "cudaEventCreate", e );
		static const float haltonx[4] = { 0.3f, 0.7f, 0.2f, 0.8f }, haltony[4] = { 0.2f, 0.8f, 0.7f, 0.3f };
		if (w->taaOn) j0 = haltonx[w->frameCycle], j1 = haltony[w->frameCycle];
		w->frameCycle = (w->frameCycle + 1) & 3;
		f.bounceCap = 1, f.roulette = 0, f.jitter = 2;
		const int cur = w->filterFlip, old = cur ^ 1;
		f.features = (uint4*)w->features.ptr, f.worldPos = (float4*)w->worldPos[cur].ptr, f.deltaDepth = (float4*)w->deltaDepth.ptr;
		FilterPlan& fp = plan.filter;
		fp.prevWorldPos = (float4*)w->worldPos[old].ptr, fp.shading = (float4*)w->shading[cur].ptr, fp.prevPixels = (float4*)w->shading[old].ptr, fp.taaOut = (float4*)w->taaOut.ptr;
		fp.moments = (float4*)w->moments[cur].ptr, fp.prevMoments = (float4*)w->moments[old].ptr;
		fp.filteredIn = (float4*)w->filtered[cur].ptr, fp.filteredOut = (float4*)w->filtered[old].ptr, fp.motion = (float2*)w->motion.ptr, fp.taa = w->taaOn ? 1u : 0u;
		fp.reprojection = MakeFilterReprojection( w->prevView, (int)f.height, j0, j1, w->prevJ0, w->prevJ1, (uint32_t)w->spp, w->filterClampDirect, w->filterClampIndirect, converge == 1 ? 0 : 1 );
		converge = 1; // (rendercore.cpp:682-688: with the filter on the accumulator starts from zero every frame)
	}
	f.probePixel = (w->probeX >= 0 && w->probeY >= 0 && w->probeX < w->width && w->probeY < w->height) ? w->probeX + w->probeY * w->width : -1;
	if ((uint64_t)f.pixelCount * (uint64_t)w->spp > PATHIDX_MASK) return fail( "Render: width * height * spp exceeds 2^28 path indices" );
	if (converge == 1) w->samplesTaken = 0;
	plan.clearAccumulator = w->samplesTaken == 0;
	// paths of this shard per sample; several samples share one pass while they fit
	const uint32_t perSample = ownTiles * TILE_PIXELS;
	uint32_t samplesPerPass = perSample ? w->pathsPerPass / perSample : 1;
	if (samplesPerPass < 1) samplesPerPass = 1;
	if (samplesPerPass > (uint32_t)w->spp) samplesPerPass = (uint32_t)w->spp;
	plan.passes = perSample ? ((uint32_t)w->spp + samplesPerPass - 1) / samplesPerPass : 0;
	// chains: the shard's tiles dealt round robin.  Automatic: two chains once a pass holds enough paths for both to fill the machine
	// (a chain's waves are persistent grids of 148 x 6 x 128 lanes; below ~4 rays per lane a wave is mostly its own tail)
	uint32_t chains = w->chains > 0 ? (uint32_t)w->chains : ((uint64_t)perSample * samplesPerPass >= (1u << 20) ? 2u : 1u);
	if (chains > ownTiles) chains = ownTiles ? ownTiles : 1;
	plan.chains = f.chainCount = chains;
	plan.maxLen = (uint32_t)w->maxPathLength, plan.bsdf = (uint32_t)w->bsdf, plan.isect = (int)isect, plan.smCount = smCount;
	// shade grid = a whole number of resident sets (the loop is grid-strided: with 8 CTAs per SM launched and 5 resident, the second,
	// partial set ran on a 60 % full machine: shade 0.40 -> 0.36 ms per C2 frame)
	static int shadeResident[4] = { 0, 0, 0, 0 };
	const int shadeKind = (w->bsdf ? 1 : 0) + (filtering ? 2 : 0);
	if (!shadeResident[shadeKind])
	{
		int n = 0;
		cudaOccupancyMaxActiveBlocksPerMultiprocessor( &n, pickShade( w->bsdf != 0, filtering ), 128, 0 );
		shadeResident[shadeKind] = n > 0 ? n : 1;
	}
	plan.shadeGrid = smCount * shadeResident[shadeKind] * B200_SHADE_GRID;
	// buffers (grown outside the enqueue: a reallocation synchronises the device)
	for (uint32_t c = 0; c < chains; c++)
	{
		Chain& ch = w->chain[c];
		const uint32_t tiles = ownTiles > c ? (ownTiles - c + chains - 1) / chains : 0;
		const uint32_t cap = tiles * TILE_PIXELS * samplesPerPass + 32;
		if (cap > ch.capacity)
		{
			for (int i = 0; ch.all( i ); i++) if ((e = ch.all( i )->reserve( (size_t)cap * 16, false, stream )) != cudaSuccess) return fail( "path buffers", e );
			ch.capacity = cap;
		}
		// filter mode: a second set of contribution slots behind the first, for the indirect light
		if (filtering && ((e = ch.slotS.reserve( (size_t)cap * 32, false, stream )) != cudaSuccess || (e = ch.slotC.reserve( (size_t)cap * 32, false, stream )) != cudaSuccess))
			return fail( "path buffers", e );
		auto& pc = plan.chain[c];
		pc.tiles = tiles, pc.capacity = tiles * TILE_PIXELS * samplesPerPass;
		for (int k = 0; k < 2; k++)
		{
			pc.rayO[k] = (float4*)ch.rayO[k].ptr, pc.rayD[k] = (float4*)ch.rayD[k].ptr, pc.state[k] = (float4*)ch.state[k].ptr, pc.info[k] = (float4*)ch.info[k].ptr;
			pc.conO[k] = (float4*)ch.conO[k].ptr, pc.conD[k] = (float4*)ch.conD[k].ptr, pc.conE[k] = (float4*)ch.conE[k].ptr;
		}
		pc.hits = (float4*)ch.hits.ptr, pc.slotS = (float4*)ch.slotS.ptr, pc.slotC = (float4*)ch.slotC.ptr;
	}
	const size_t counterWords = (size_t)COUNTER_STRIDE * chains * (plan.passes ? plan.passes : 1);
	if ((e = w->counters.reserve( counterWords * 4, false, stream )) != cudaSuccess || (e = w->passParams.reserve( sizeof( PassParams ) * (plan.passes + 1), false, stream )) != cudaSuccess ||
		(e = w->hostCounters.reserve( counterWords * 4 + 64 )) != cudaSuccess || (e = w->hostPass.reserve( sizeof( PassParams ) * (plan.passes + 1) )) != cudaSuccess)
		return fail( "frame buffers", e );
	plan.trace = baseArgs();
	plan.shadePool = (const float4*)shadePool.ptr, plan.vnormPool = (const float4*)vnormPool.ptr, plan.instShadePool = (const float4*)instShadePool.ptr;
	plan.materials = (const float4*)w->materials.ptr;
	plan.accumulator = w->accumulator(), plan.frame = (float4*)w->frame.ptr;
	plan.passDev = (PassParams*)w->passParams.ptr, plan.passHost = (PassParams*)w->hostPass.ptr;
	plan.countersDev = (uint32_t*)w->counters.ptr, plan.countersHost = (uint32_t*)w->hostCounters.ptr, plan.fetchCounters = (uint32_t*)fetchCounter.ptr;
	// the gather is part of the frame when the peers are open and the frame is sharded over exactly them (Setting gatherInRender = 0: the
	// host calls b200_gather_finalize itself)
	plan.gather = (w->gatherInRender && w->peerWorld > 1 && w->shardCount == w->peerWorld && w->shardIndex == w->peerRank && !w->accumExternal) ? 1u : 0u;
	if (plan.gather)
	{
		plan.peerWorld = w->peerWorld, plan.peerRank = w->peerRank, plan.peers = w->peers;
		plan.peerEpochWord = (uint32_t*)w->peerFlags.ptr + MAX_PEERS + 1, plan.peerTimeout = (uint32_t*)w->peerFlags.ptr + MAX_PEERS;
		plan.peerTimeoutHost = (uint32_t*)((char*)w->hostCounters.ptr + counterWords * 4);
		plan.peerLimitTicks = peerLimitTicks();
		*plan.peerTimeoutHost = 0;
	}
	// per-wave timers: every pass while that stays a few hundred events, else the first pass only (its times are scaled to the frame's
	// ray counts below)
	plan.timedPasses = !w->waveTimers ? 0 : (plan.passes * chains * plan.maxLen <= 96 ? plan.passes : 1);
	const size_t timersNeeded = (size_t)plan.timedPasses * chains * plan.maxLen * 5;
	while (w->timers.size() < timersNeeded)
	{
		cudaEvent_t ev;
		if ((e = cudaEventCreate( &ev )) != cudaSuccess) return fail( "cudaEventCreate", e );
		w->timers.push_back( ev );
	}
	// this frame's PassParams
	for (uint32_t pass = 0; pass < plan.passes; pass++)
	{
		PassParams& P = plan.passHost[pass];
		memset( &P, 0, sizeof( P ) );
		for (int a = 0; a < 3; a++) P.pos[a] = vp[a], P.p1[a] = vp[3 + a], P.right[a] = vp[6 + a] - vp[3 + a], P.up[a] = vp[9 + a] - vp[3 + a];
		P.aperture = vp[12], P.spreadAngle = vp[13], P.distortion = vp[16];
		if (filtering)
		{
			// jitteredView: p1 + (j0 / width) right + (j1 / height) up; no lens, no distortion (rendercore.cpp:738-746)
			for (int a = 0; a < 3; a++) P.p1[a] = vp[3 + a] + ((j0 / (float)f.width) * P.right[a] + (j1 / (float)f.height) * P.up[a]);
			P.aperture = 0, P.distortion = 0;
		}
		const uint32_t s0 = pass * samplesPerPass;
		P.samplesThisPass = (uint32_t)w->spp - s0 < samplesPerPass ? (uint32_t)w->spp - s0 : samplesPerPass;
		// pathIdx = pixel + sample * pixelCount must stay below 2^28 (the path word keeps 4 flag bits): the sample index wraps after `period` samples and the wrap
		// count goes into the salt instead ("samplesTaken * 7907 + pathLength * 91771" in the reference's shade call)
		const uint32_t globalSample = (uint32_t)w->sampleOffset + w->samplesTaken + s0;
		const uint32_t period = PATHIDX_MASK / f.pixelCount - (uint32_t)w->spp + 1;
		P.sampleBase = globalSample % period;
		P.blue = (w->haveBlueNoise && globalSample < period) ? 1u : 0u; // not after the sample index wrapped
		P.R0 = (globalSample / period + 1) * 7907u;
		P.finalScale = 1.0f / (float)(w->samplesTaken + (uint32_t)w->spp);
		if (filtering)
		{
			// Every filtered frame starts from sample 0 with the same seed (rendercore.cpp:687 "same seed means same noise"): in the reference the
			// frames differ through the rotating blue-noise slot alone.  A host that has supplied the tables gets exactly that; without
			// them the salt changes with the frame instead, or the temporal filter would average a still noise pattern.
			P.blueSlot = w->blueSlot;
			if (!w->haveBlueNoise) P.R0 = (w->filterFrames + 1) * 7907u;
		}
	}
	memset( &stats.totalRays, 0, (char*)&stats.probedInstid - (char*)&stats.totalRays );
	plan.l2Persist = l2Persist;
	if (l2Persist != w->l2PersistApplied || plan.trace.nodes != w->l2WindowNodes)
	{
		// (the streams the wave kernels of a frame run on; part of the plan, so a change re-captures the graph)
		for (Chain& c : w->chain) if (c.main && (applyL2Window( c.main ) || applyL2Window( c.connect ))) return 1;
		w->l2PersistApplied = l2Persist, w->l2WindowNodes = plan.trace.nodes;
	}
	if (plan.passes)
	{
		// A frame whose plan equals the previous frame's is replayed as a graph (captured on its second appearance, so that one-off
		// frames never pay for a capture); anything else is enqueued on the streams directly
		std::vector<uint8_t> key( (const uint8_t*)&plan, (const uint8_t*)&plan + sizeof( plan ) );
		uint32_t launches = 0;
		bool launched = false;
		if (w->useGraph && w->graphExec && key == w->graphKey)
		{
			if ((e = cudaGraphLaunch( w->graphExec, stream )) != cudaSuccess) return fail( "frame graph launch", e );
			launched = true, w->graphLaunches++;
		}
		else if (w->useGraph && key == w->lastKey)
		{
			cudaGraph_t graph = nullptr;
			if (w->graphExec) { cudaGraphExecDestroy( w->graphExec ); w->graphExec = nullptr; }
			cudaStreamSynchronize( stream ); // work the caller queued before this frame is not part of the capture
			if ((e = cudaStreamBeginCapture( w->captureStream, cudaStreamCaptureModeThreadLocal )) != cudaSuccess) return fail( "frame graph capture", e );
			e = enqueueFrame( *w, plan, w->captureStream, true, &launches );
			cudaError_t e2 = cudaStreamEndCapture( w->captureStream, &graph );
			if (e == cudaSuccess) e = e2;
			if (e == cudaSuccess) e = cudaGraphInstantiate( &w->graphExec, graph, 0 );
			if (graph) cudaGraphDestroy( graph );
			if (e != cudaSuccess) { w->graphExec = nullptr; cudaGetLastError(); return fail( "frame graph capture", e ); }
			w->graphKey = key, w->graphCaptures++;
			if ((e = cudaGraphLaunch( w->graphExec, stream )) != cudaSuccess) return fail( "frame graph launch", e );
			launched = true, w->graphLaunches++;
		}
		if (!launched && (e = enqueueFrame( *w, plan, stream, false, &launches )) != cudaSuccess) return fail( "Render", e );
		w->lastKey.swap( key );
		launches = 0; // kernels of the frame: per pass and chain one generate, per path length extend + shade (+ connect), one accumulate
		for (uint32_t c = 0; c < chains; c++) if (plan.chain[c].tiles) launches += plan.passes * (2 + plan.maxLen * (2 + (plan.anyLight ? 1 : 0)));
		if (plan.gather) launches += 3;
		stats.kernelLaunches += launches;
		stats.hostEnqueueMs = hostMs();
		if ((e = cudaStreamSynchronize( stream )) != cudaSuccess) return fail( "Render", e );
		stats.hostWaitMs = hostMs() - stats.hostEnqueueMs;
		cudaEventElapsedTime( &stats.renderTime, w->frameBegin, w->frameEnd );
		stats.filterTime = 0;
		if (filtering)
		{
			// what this frame leaves for the next one (rendercore.cpp:855-861, :878): the blended image as TAA history, world positions, moments and
			// the first reconstruction pass change places with their predecessors
			cudaEventElapsedTime( &stats.filterTime, w->filterBegin, w->filterEnd );
			if (w->taaOn) std::swap( w->taaOut, w->shading[w->filterFlip] );
			w->filterFlip ^= 1;
			w->prevJ0 = j0, w->prevJ1 = j1;
			w->blueSlot = (w->blueSlot + 1) & 255u, w->filterFrames++;
			memcpy( w->prevView, vp, 48 );
		}
		stats.gatherMs = stats.gatherKernelMs = 0;
		if (plan.gather)
		{
			cudaEventElapsedTime( &stats.gatherMs, w->gatherWait, w->frameEnd ), cudaEventElapsedTime( &stats.gatherKernelMs, w->gatherBegin, w->gatherEnd );
			if (*plan.peerTimeoutHost)
			{
				cudaMemsetAsync( plan.peerTimeout, 0, 4, stream );
				return fail( "Render: a peer did not arrive within the time limit (Setting peerTimeoutMs); the frame was not gathered" );
			}
		}
		// statistics: ray counts of every pass, wave times of the timed ones
		double timedRays[3] = { 0, 0, 0 }, allRays[3] = { 0, 0, 0 }, timedShadow = 0, allShadow = 0, timedPaths = 0, allPaths = 0; // classes: primary, first bounce, deeper
		float classMs[3] = { 0, 0, 0 }, shadowMs = 0, shadeMs = 0;
		for (uint32_t pass = 0; pass < plan.passes; pass++) for (uint32_t c = 0; c < chains; c++)
		{
			const uint32_t* hc = plan.countersHost + (size_t)(pass * chains + c) * COUNTER_STRIDE;
			const bool timed = pass < plan.timedPasses;
			for (uint32_t len = 1; len <= plan.maxLen; len++)
			{
				const uint32_t nExt = hc[extCounter( (int)len )], nShadow = hc[shadowCounter( (int)len )];
				const int cls = len == 1 ? 0 : len == 2 ? 1 : 2;
				allRays[cls] += nExt, allShadow += nShadow, allPaths += nExt;
				if (len == 1) stats.primaryRayCount += nExt; else if (len == 2) stats.bounce1RayCount += nExt; else stats.deepRayCount += nExt;
				stats.totalExtensionRays += nExt, stats.totalShadowRays += nShadow;
				if (timed && plan.chain[c].tiles)
				{
					const cudaEvent_t* t = &w->timers[((size_t)(pass * chains + c) * plan.maxLen + (len - 1)) * 5];
					float tExt = 0, tShade = 0, tCon = 0;
					cudaEventElapsedTime( &tExt, t[0], t[1] ), cudaEventElapsedTime( &tShade, t[1], t[2] );
					if (plan.anyLight) cudaEventElapsedTime( &tCon, t[3], t[4] );
					classMs[cls] += tExt, shadeMs += tShade, shadowMs += tCon;
					timedRays[cls] += nExt, timedShadow += nShadow, timedPaths += nExt;
				}
			}
			if (f.probePixel >= 0 && pass == 0 && hc[PROBE_BASE + 3])
			{
				stats.probedInstid = (int32_t)hc[PROBE_BASE], stats.probedTriid = (int32_t)hc[PROBE_BASE + 1];
				memcpy( &stats.probedDist, &hc[PROBE_BASE + 2], 4 );
			}
		}
		// (chains and connect waves overlap: a wave's time is its own begin-to-end span, so the class times add up to more than the frame)
		auto scaled = []( float ms, double timed, double all ) { return timed > 0 ? (float)(ms * all / timed) : 0.f; };
		stats.traceTime0 = scaled( classMs[0], timedRays[0], allRays[0] ), stats.traceTime1 = scaled( classMs[1], timedRays[1], allRays[1] );
		stats.traceTimeX = scaled( classMs[2], timedRays[2], allRays[2] );
		stats.shadowTraceTime = scaled( shadowMs, timedShadow, allShadow ), stats.shadeTime = scaled( shadeMs, timedPaths, allPaths );
	}
	stats.hostTotalMs = hostMs();
	w->samplesTaken += (uint32_t)w->spp;
	stats.frameGraphLaunches = w->graphLaunches, stats.frameChains = plan.chains, stats.framePasses = plan.passes;
	stats.totalRays = stats.totalExtensionRays + stats.totalShadowRays;
	w->frameCounter++;
	return 0;
}

int RenderCore::Finalize( uint32_t totalSamples )
{
	WavefrontState* w = wf;
	if (!w || w->width == 0) return fail( "Finalize: SetTarget has not been called" );
	if (!totalSamples) return fail( "Finalize: zero samples" );
	const uint32_t px = (uint32_t)w->width * (uint32_t)w->height;
	finalizeKernel<<<(px + 255) / 256, 256, 0, stream>>>( w->accumulator(), (float4*)w->frame.ptr, px, 1.0f / (float)totalSamples );
	stats.kernelLaunches++;
	cudaError_t e = cudaStreamSynchronize( stream );
	return e == cudaSuccess ? 0 : fail( "Finalize", e );
}

// ---- peer-memory gather: PeerHandle -> (exchange the 128-byte blobs between the ranks) -> OpenPeers -> GatherFinalize per frame ----
// out128: the CUDA IPC handle of this core's own accumulator, then the one of its arrival flags.  The accumulator must be the
// core's own (cudaMalloc'ed) one and must not be reallocated afterwards: call after SetTarget.
int RenderCore::PeerHandle( void* out128 )
{
	WavefrontState* w = wf;
	if (!w || w->width == 0) return fail( "PeerHandle: SetTarget has not been called" );
	if (w->accumExternal) return fail( "PeerHandle: the accumulator is caller-owned (b200_set_accumulator_buffer); peers map the core's own one" );
	if (!out128) return fail( "PeerHandle: null output" );
	cudaError_t e;
	// MAX_PEERS arrival words (written by the peers), the timeout flag, this rank's epoch word
	if ((e = w->peerFlags.reserve( (MAX_PEERS + 2) * 4, false, stream )) != cudaSuccess) return fail( "peer flags", e );
	if ((e = cudaMemsetAsync( w->peerFlags.ptr, 0, (MAX_PEERS + 2) * 4, stream )) != cudaSuccess) return fail( "peer flags", e );
	if ((e = cudaStreamSynchronize( stream )) != cudaSuccess) return fail( "peer flags", e );
	w->dropGraph();
	cudaIpcMemHandle_t h[2];
	static_assert( sizeof( cudaIpcMemHandle_t ) == 64, "IPC handle size" );
	if ((e = cudaIpcGetMemHandle( &h[0], w->accumOwn.ptr )) != cudaSuccess) return fail( "cudaIpcGetMemHandle(accumulator)", e );
	if ((e = cudaIpcGetMemHandle( &h[1], w->peerFlags.ptr )) != cudaSuccess) return fail( "cudaIpcGetMemHandle(flags)", e );
	memcpy( out128, h, 128 );
	return 0;
}

// handles: world x 128 bytes, entry r = what rank r's PeerHandle returned (entry `rank` is not opened: it is this core)
int RenderCore::OpenPeers( const void* handles, int world, int rank )
{
	WavefrontState* w = wf;
	if (!w || w->width == 0 || !w->peerFlags.ptr) return fail( "OpenPeers: call PeerHandle first" );
	if (!handles || world < 1 || world > MAX_PEERS || rank < 0 || rank >= world) return fail( "OpenPeers: bad arguments" );
	w->closePeers();
	for (int r = 0; r < world; r++)
	{
		if (r == rank) { w->peers.acc[r] = (float4*)w->accumOwn.ptr, w->peers.flags[r] = (uint32_t*)w->peerFlags.ptr; continue; }
		cudaIpcMemHandle_t h[2];
		memcpy( h, (const char*)handles + (size_t)r * 128, 128 );
		for (int k = 0; k < 2; k++)
		{
			void* p = nullptr;
			cudaError_t e = cudaIpcOpenMemHandle( &p, h[k], cudaIpcMemLazyEnablePeerAccess );
			if (e != cudaSuccess) { w->closePeers(); return fail( "cudaIpcOpenMemHandle", e ); }
			w->peerMapped[2 * r + k] = p;
		}
		w->peers.acc[r] = (float4*)w->peerMapped[2 * r], w->peers.flags[r] = (uint32_t*)w->peerMapped[2 * r + 1];
	}
	w->peerWorld = world, w->peerRank = rank;
	return 0;
}

// the peer barrier's patience in clock64 ticks (Setting peerTimeoutMs at the SM clock)
long long RenderCore::peerLimitTicks() const
{
	static int clockKHz = 0; // (asked once: this attribute goes down to the kernel driver and costs about a millisecond)
	if (!clockKHz) cudaDeviceGetAttribute( &clockKHz, cudaDevAttrClockRate, device );
	return (long long)((double)(wf ? wf->peerTimeoutMs : 2000.f) * (clockKHz > 0 ? clockKHz : 1965000));
}

int RenderCore::ClosePeers()
{
	if (wf) { cudaStreamSynchronize( stream ); wf->closePeers(); }
	return 0;
}

// One frame's gather: arrive (everyone has rendered), pull + finalize, arrive again (everyone has pulled: the accumulators may
// be written again).  After it the local accumulator holds the whole image, like after the all-reduce it replaces.
int RenderCore::GatherFinalize( uint32_t totalSamples )
{
	WavefrontState* w = wf;
	if (!w || w->width == 0) return fail( "GatherFinalize: SetTarget has not been called" );
	if (w->peerWorld < 1) return fail( "GatherFinalize: OpenPeers has not been called" );
	if (!totalSamples) return fail( "GatherFinalize: zero samples" );
	if (w->shardCount != w->peerWorld || w->shardIndex != w->peerRank) return fail( "GatherFinalize: shardCount / shardIndex differ from the peer table" );
	uint32_t* timeout = (uint32_t*)w->peerFlags.ptr + MAX_PEERS;
	const uint32_t tilesX = ((uint32_t)w->width + TILE_W - 1) / TILE_W;
	const long long limitTicks = peerLimitTicks();
	uint32_t* epochWord = (uint32_t*)w->peerFlags.ptr + MAX_PEERS + 1;
	cudaError_t e;
	peerBarrierKernel<<<1, 32, 0, stream>>>( w->peers, w->peerWorld, w->peerRank, epochWord, timeout, limitTicks );
	cudaEventRecord( w->gatherBegin, stream ); // the pull kernel alone (everyone has arrived): what the NVLink figure of bench.py divides by
	gatherFinalizeKernel<<<smCount * 8, 256, 0, stream>>>( w->peers, w->peerWorld, w->peerRank, (float4*)w->frame.ptr, (uint32_t)w->width, (uint32_t)w->height, tilesX,
		tileSkewFor( (uint32_t)w->peerWorld ), 1.0f / (float)totalSamples, nullptr, timeout );
	cudaEventRecord( w->gatherEnd, stream );
	peerBarrierKernel<<<1, 32, 0, stream>>>( w->peers, w->peerWorld, w->peerRank, epochWord, timeout, limitTicks );
	stats.kernelLaunches += 3;
	uint32_t timedOut = 0;
	e = cudaMemcpyAsync( &timedOut, timeout, 4, cudaMemcpyDeviceToHost, stream );
	if (e == cudaSuccess) e = cudaStreamSynchronize( stream );
	if (e != cudaSuccess) return fail( "GatherFinalize", e );
	cudaEventElapsedTime( &stats.gatherKernelMs, w->gatherBegin, w->gatherEnd );
	if (timedOut)
	{
		// this frame's gather did not happen (the pull kernel saw the flag and left); the next one starts with a clean flag
		cudaMemsetAsync( timeout, 0, 4, stream );
		return fail( "GatherFinalize: a peer did not arrive within the time limit (Setting peerTimeoutMs); the frame was not gathered" );
	}
	return 0;
}

int RenderCore::ReadAccumulator( float* dst )
{
	WavefrontState* w = wf;
	if (!w || w->width == 0) return fail( "ReadAccumulator: SetTarget has not been called" );
	cudaError_t e = cudaMemcpyAsync( dst, w->accumulator(), (size_t)w->width * w->height * 16, cudaMemcpyDeviceToHost, stream );
	if (e == cudaSuccess) e = cudaStreamSynchronize( stream );
	return e == cudaSuccess ? 0 : fail( "ReadAccumulator", e );
}

int RenderCore::ReadFrame( float* dst )
{
	WavefrontState* w = wf;
	if (!w || w->width == 0) return fail( "ReadFrame: SetTarget has not been called" );
	cudaError_t e = cudaMemcpyAsync( dst, w->frame.ptr, (size_t)w->width * w->height * 16, cudaMemcpyDeviceToHost, stream );
	if (e == cudaSuccess) e = cudaStreamSynchronize( stream );
	return e == cudaSuccess ? 0 : fail( "ReadFrame", e );
}

// The same without making the caller wait: the finished frame is copied aside on the core's stream (0.05 ms at 4K) and leaves for the
// host on a stream of its own, so that the next Render() runs beside the 132 MB of a 4K frame on the bus.  WaitFrame() returns when the
// last such frame has arrived; a second ReadFrameAsync before that is fine (its device copy waits for the first read to be through).
int RenderCore::ReadFrameAsync( float* dst )
{
	WavefrontState* w = wf;
	if (!w || w->width == 0) return fail( "ReadFrameAsync: SetTarget has not been called" );
	if (!dst) return fail( "ReadFrameAsync: null destination" );
	const size_t bytes = (size_t)w->width * w->height * 16;
	cudaError_t e;
	if (!w->readStream)
	{
		if ((e = cudaStreamCreateWithFlags( &w->readStream, cudaStreamNonBlocking )) != cudaSuccess) return fail( "ReadFrameAsync", e );
		if ((e = cudaEventCreateWithFlags( &w->frameCopied, cudaEventDisableTiming )) != cudaSuccess) return fail( "ReadFrameAsync", e );
		if ((e = cudaEventCreateWithFlags( &w->frameRead, cudaEventDisableTiming )) != cudaSuccess) return fail( "ReadFrameAsync", e );
	}
	if (w->frameStage.bytes < bytes)
	{
		if (w->framePending && (e = cudaEventSynchronize( w->frameRead )) != cudaSuccess) return fail( "ReadFrameAsync", e );
		if ((e = w->frameStage.reserve( bytes, false, stream )) != cudaSuccess) return fail( "ReadFrameAsync staging", e );
	}
	if (w->framePending && (e = cudaStreamWaitEvent( stream, w->frameRead, 0 )) != cudaSuccess) return fail( "ReadFrameAsync", e );
	if ((e = cudaMemcpyAsync( w->frameStage.ptr, w->frame.ptr, bytes, cudaMemcpyDeviceToDevice, stream )) != cudaSuccess) return fail( "ReadFrameAsync", e );
	if ((e = cudaEventRecord( w->frameCopied, stream )) != cudaSuccess) return fail( "ReadFrameAsync", e );
	if ((e = cudaStreamWaitEvent( w->readStream, w->frameCopied, 0 )) != cudaSuccess) return fail( "ReadFrameAsync", e );
	if ((e = cudaMemcpyAsync( dst, w->frameStage.ptr, bytes, cudaMemcpyDeviceToHost, w->readStream )) != cudaSuccess) return fail( "ReadFrameAsync", e );
	if ((e = cudaEventRecord( w->frameRead, w->readStream )) != cudaSuccess) return fail( "ReadFrameAsync", e );
	w->framePending = true;
	return 0;
}

int RenderCore::WaitFrame()
{
	WavefrontState* w = wf;
	if (!w || !w->framePending) return 0;
	const cudaError_t e = cudaEventSynchronize( w->frameRead );
	w->framePending = false;
	return e == cudaSuccess ? 0 : fail( "WaitFrame", e );
}

// The per-pixel buffers of the last filtered frame (feature outputs for a host that wants them, and the tests): 0 features (uint4), 1 world
// positions, 2 depth derivatives, 3 the indirect accumulator plane, 4 motion (float2), 5 luminance moments, 6 the first reconstruction pass
// (packed light, the next frame's history), 7 the TAA history image.  16 bytes per pixel but for the motion vectors (8).
int RenderCore::ReadFilterBuffer( int which, void* dst )
{
	WavefrontState* w = wf;
	if (!w || w->width == 0 || !w->filterPixels) return fail( "ReadFilterBuffer: no filtered frame has been rendered" );
	if (!dst) return fail( "ReadFilterBuffer: null destination" );
	const size_t px = (size_t)w->width * w->height;
	const int last = w->filterFlip ^ 1; // Render flipped the pairs when the frame was done
	const void* src = nullptr;
	size_t bytes = px * 16;
	switch (which)
	{
	case 0: src = w->features.ptr; break;
	case 1: src = w->worldPos[last].ptr; break;
	case 2: src = w->deltaDepth.ptr; break;
	case 3: src = (const char*)w->accumOwn.ptr + px * 16; break;
	case 4: src = w->motion.ptr, bytes = px * 8; break;
	case 5: src = w->moments[last].ptr; break;
	case 6: src = w->filtered[w->filterFlip].ptr; break; // (this frame's "filteredOut" is the next frame's "filteredIn")
	case 7: src = w->shading[last].ptr; break;
	default: return fail( "ReadFilterBuffer: buffers 0..7" );
	}
	cudaError_t e = cudaMemcpyAsync( dst, src, bytes, cudaMemcpyDeviceToHost, stream );
	if (e == cudaSuccess) e = cudaStreamSynchronize( stream );
	return e == cudaSuccess ? 0 : fail( "ReadFilterBuffer", e );
}

void* RenderCore::DeviceAccumulator() { return wf ? (void*)wf->accumulator() : nullptr; }

} // namespace b200
