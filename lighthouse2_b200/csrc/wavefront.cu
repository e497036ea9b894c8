// wavefront.cu -- the frame around the two waves: ray generation, minimal diffuse shade, fused connect, finalize.
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
//   * the shade stage is the minimal diffuse model SURVEY.md 7 step 6 asks for (Lambert, one point light with NEE,
//     constant sky, fixed path length); the full LH2 material system is a "next" row
// Pixels can be sharded over several cores/GPUs by interleaved 32x8 tiles (Setting shardIndex/shardCount) or by sample
// (Setting sampleOffset); per-path random streams depend only on (pixel, sample, bounce), so any sharding gives the same
// image and the accumulators of all shards simply add up.
#include "rendercore.h"
#include <cstring>

namespace b200 {

namespace {

constexpr int TILE_W = 32, TILE_H = 8, TILE_PIXELS = TILE_W * TILE_H;
constexpr float INVPI = 0.31830988618379067154f, TWOPI = 6.28318530717958647692f;

struct WaveCounters { uint32_t extensionRays, shadowRays, pad0, pad1; }; // one per path length (index 0: probe result)

struct FrameParams {
	float pos[3], p1[3], right[3], up[3]; // right = p2 - p1, up = p3 - p1
	float lightPos[3], lightRadiance[3], sky[3];
	float epsilon;
	uint32_t width, height, pixelCount;
	uint32_t shardIndex, shardCount, ownTiles, tilesX, tilesY;
	uint32_t sampleBase, samplesThisPass; // global index of the first sample of this pass
	uint32_t R0;                          // per-frame RNG salt
	uint32_t hasLight, jitter, materialCount;
	int32_t probePixel;
};

// lib/CUDA/shared_kernel_code/tools_shared.h:60-62
__device__ __forceinline__ uint32_t WangHash( uint32_t s ) { s = (s ^ 61) ^ (s >> 16), s *= 9, s = s ^ (s >> 4), s *= 0x27d4eb2d, s = s ^ (s >> 15); return s; }
__device__ __forceinline__ uint32_t RandomInt( uint32_t& s ) { s ^= s << 13, s ^= s >> 17, s ^= s << 5; return s; }
__device__ __forceinline__ float RandomFloat( uint32_t& s ) { return RandomInt( s ) * 2.3283064365387e-10f; }

__device__ __forceinline__ uint32_t laneMaskLt()
{
	uint32_t m;
	asm( "mov.u32 %0, %%lanemask_lt;" : "=r"( m ) );
	return m;
}

// warp-aggregated append (stream compaction): all 32 lanes call; returns the slot of lanes with want == true
__device__ __forceinline__ uint32_t warpAppend( uint32_t* counter, bool want )
{
	const uint32_t m = __ballot_sync( 0xffffffffu, want );
	if (!m) return 0;
	const uint32_t leader = __ffs( m ) - 1;
	uint32_t base = 0;
	if ((threadIdx.x & 31) == leader) base = atomicAdd( counter, __popc( m ) );
	base = __shfl_sync( 0xffffffffu, base, leader );
	return base + __popc( m & laneMaskLt() );
}

// ---- primary rays (GPU cores: rendercore_optix7/optix/.optix.cu:66-123, RenderCore_OptixPrime_B/kernels/camera.h:40-94) ----------
// u = (x + r0) / w, v = (y + r1) / h, P = p1 + u right + v up, D = normalize(P - pos); pinhole (aperture is not sampled).
__global__ void __launch_bounds__( 256 ) generateEyeRaysKernel( const FrameParams f, float4* __restrict__ rayO, float4* __restrict__ rayD,
	float4* __restrict__ pathState, WaveCounters* counters )
{
	const uint32_t total = f.ownTiles * TILE_PIXELS * f.samplesThisPass;
	const uint32_t stride = gridDim.x * blockDim.x;
	for (uint32_t base = blockIdx.x * blockDim.x + (threadIdx.x & ~31u); base < total; base += stride)
	{
		const uint32_t job = base + (threadIdx.x & 31);
		bool valid = job < total;
		uint32_t x = 0, y = 0, sample = 0;
		if (valid)
		{
			const uint32_t perSample = f.ownTiles * TILE_PIXELS;
			sample = job / perSample;
			const uint32_t q = job - sample * perSample, localTile = q / TILE_PIXELS, inTile = q % TILE_PIXELS;
			const uint32_t tile = localTile * f.shardCount + f.shardIndex;
			x = (tile % f.tilesX) * TILE_W + (inTile % TILE_W), y = (tile / f.tilesX) * TILE_H + inTile / TILE_W;
			valid = x < f.width && y < f.height;
		}
		const uint32_t slot = warpAppend( &counters[1].extensionRays, valid );
		if (!valid) continue;
		const uint32_t pixelIdx = x + y * f.width, pathIdx = pixelIdx + (f.sampleBase + sample) * f.pixelCount;
		float r0 = 0.5f, r1 = 0.5f;
		if (f.jitter)
		{
			uint32_t seed = WangHash( pathIdx + f.R0 );
			r0 = RandomFloat( seed ), r1 = RandomFloat( seed );
		}
		const float u = ((float)x + r0) / (float)f.width, v = ((float)y + r1) / (float)f.height;
		const float dx = f.p1[0] + u * f.right[0] + v * f.up[0] - f.pos[0];
		const float dy = f.p1[1] + u * f.right[1] + v * f.up[1] - f.pos[1];
		const float dz = f.p1[2] + u * f.right[2] + v * f.up[2] - f.pos[2];
		const float il = 1.0f / sqrtf( dx * dx + dy * dy + dz * dz );
		rayO[slot] = make_float4( f.pos[0], f.pos[1], f.pos[2], f.epsilon );
		rayD[slot] = make_float4( dx * il, dy * il, dz * il, 1e34f );
		pathState[slot] = make_float4( 1.f, 1.f, 1.f, __uint_as_float( pathIdx ) );
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

// ---- shade (model: lib/RenderCore_PrimeRef/kernels/pathtracer.h:46-165 reduced to Lambert + one point light) ----------------------
__global__ void __launch_bounds__( 128 ) shadeKernel( const FrameParams f, const int pathLength, const int maxPathLength,
	const float4* __restrict__ rayO, const float4* __restrict__ rayD, const float4* __restrict__ pathState, const float4* __restrict__ hits,
	float4* __restrict__ rayOout, float4* __restrict__ rayDout, float4* __restrict__ pathStateOut,
	float4* __restrict__ conO, float4* __restrict__ conD, float4* __restrict__ conE,
	const float4* __restrict__ shade, const float4* __restrict__ instShade, const float4* __restrict__ materials,
	float4* accumulator, WaveCounters* counters )
{
	const uint32_t count = counters[pathLength].extensionRays;
	const uint32_t stride = gridDim.x * blockDim.x;
	for (uint32_t base = blockIdx.x * blockDim.x + (threadIdx.x & ~31u); base < count; base += stride)
	{
		const uint32_t job = base + (threadIdx.x & 31);
		bool wantShadow = false, wantExt = false;
		float sox = 0, soy = 0, soz = 0, lx = 0, ly = 0, lz = 0, ldist = 0, ex = 0, ey = 0, ez = 0;
		float eox = 0, eoy = 0, eoz = 0, rx = 0, ry = 0, rz = 0, tx = 0, ty = 0, tz = 0;
		uint32_t pathIdx = 0, pixelIdx = 0;
		if (job < count)
		{
			const float4 O4 = __ldg( rayO + job ), D4 = __ldg( rayD + job ), T4 = __ldg( pathState + job ), hit = __ldg( hits + job );
			pathIdx = __float_as_uint( T4.w ), pixelIdx = pathIdx % f.pixelCount;
			const int prim = (int)__float_as_uint( hit.z );
			float* px = reinterpret_cast<float*>( accumulator + pixelIdx );
			if (prim < 0)
			{
				// sky (the reference samples the skydome, pathtracer.h:77-82; this core keeps a constant sky colour)
				atomicAdd( px, T4.x * f.sky[0] ), atomicAdd( px + 1, T4.y * f.sky[1] ), atomicAdd( px + 2, T4.z * f.sky[2] );
			}
			else
			{
				const uint32_t inst = __float_as_uint( hit.y );
				if (pathLength == 1 && (int)pixelIdx == f.probePixel && pathIdx < f.pixelCount)
					counters[0].extensionRays = inst, counters[0].shadowRays = (uint32_t)prim, counters[0].pad0 = __float_as_uint( hit.w ); // pathtracer.h:94-97
				const float4 m0 = __ldg( instShade + inst * 4 ), m1 = __ldg( instShade + inst * 4 + 1 ), m2 = __ldg( instShade + inst * 4 + 2 ),
							 info = __ldg( instShade + inst * 4 + 3 );
				const float4 sh = __ldg( shade + __float_as_uint( info.x ) + prim );
				// world-space geometric normal: normalize( transform.TransformVector( N ) ) (bvh.cpp:616-619)
				float nx = m0.x * sh.x + m0.y * sh.y + m0.z * sh.z, ny = m1.x * sh.x + m1.y * sh.y + m1.z * sh.z, nz = m2.x * sh.x + m2.y * sh.y + m2.z * sh.z;
				const float nl = 1.0f / sqrtf( nx * nx + ny * ny + nz * nz );
				nx *= nl, ny *= nl, nz *= nl;
				if (D4.x * nx + D4.y * ny + D4.z * nz > 0) nx = -nx, ny = -ny, nz = -nz; // faceDir (pathtracer.h:117-118)
				const uint32_t mat = __float_as_uint( sh.w );
				float cr = 0.8f, cg = 0.8f, cb = 0.8f;
				if (mat < f.materialCount) { const float4 c = __ldg( materials + mat ); cr = c.x, cg = c.y, cb = c.z; }
				const float ix = O4.x + hit.w * D4.x, iy = O4.y + hit.w * D4.y, iz = O4.z + hit.w * D4.z;
				// next event estimation towards the point light (pathtracer.h:121-145)
				if (f.hasLight)
				{
					float Lx = f.lightPos[0] - ix, Ly = f.lightPos[1] - iy, Lz = f.lightPos[2] - iz;
					const float d2 = Lx * Lx + Ly * Ly + Lz * Lz, dist = sqrtf( d2 ), id = 1.0f / dist;
					Lx *= id, Ly *= id, Lz *= id;
					const float NdotL = Lx * nx + Ly * ny + Lz * nz;
					if (NdotL > 0 && dist > 4 * f.epsilon)
					{
						const float s = NdotL * INVPI / d2;
						ex = T4.x * cr * f.lightRadiance[0] * s, ey = T4.y * cg * f.lightRadiance[1] * s, ez = T4.z * cb * f.lightRadiance[2] * s;
						SafeOrigin( ix, iy, iz, Lx, Ly, Lz, nx, ny, nz, f.epsilon, sox, soy, soz );
						lx = Lx, ly = Ly, lz = Lz, ldist = dist - 2 * f.epsilon;
						wantShadow = true;
					}
				}
				// diffuse bounce: cosine-weighted direction, Lambert throughput = albedo (bsdf * cos / pdf)
				if (pathLength < maxPathLength)
				{
					uint32_t seed = WangHash( pathIdx * 17 + f.R0 + (uint32_t)pathLength * 91771u ); // pathtracer.h:155; rendercore.cpp (R0 per bounce)
					const float r0 = RandomFloat( seed ), r1 = RandomFloat( seed );
					const float term1 = TWOPI * r0, term2 = sqrtf( 1 - r1 );
					float sn, cs;
					__sincosf( term1, &sn, &cs );
					Tangent2World( cs * term2, sn * term2, sqrtf( r1 ), nx, ny, nz, rx, ry, rz );
					const float rl = 1.0f / sqrtf( rx * rx + ry * ry + rz * rz );
					rx *= rl, ry *= rl, rz *= rl;
					SafeOrigin( ix, iy, iz, rx, ry, rz, nx, ny, nz, f.epsilon, eox, eoy, eoz );
					tx = T4.x * cr, ty = T4.y * cg, tz = T4.z * cb;
					wantExt = true;
				}
			}
		}
		const uint32_t sslot = warpAppend( &counters[pathLength].shadowRays, wantShadow );
		if (wantShadow)
		{
			conO[sslot] = make_float4( sox, soy, soz, f.epsilon );
			conD[sslot] = make_float4( lx, ly, lz, ldist );
			conE[sslot] = make_float4( ex, ey, ez, __uint_as_float( pixelIdx ) );
		}
		const uint32_t eslot = warpAppend( &counters[pathLength + 1].extensionRays, wantExt );
		if (wantExt)
		{
			rayOout[eslot] = make_float4( eox, eoy, eoz, f.epsilon );
			rayDout[eslot] = make_float4( rx, ry, rz, 1e34f );
			pathStateOut[eslot] = make_float4( tx, ty, tz, __uint_as_float( pathIdx ) );
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

constexpr int MAX_PATH_LENGTH = 16; // MAXPATHLENGTH of the optix7 core (core_settings.h:24)

} // namespace

struct WavefrontState {
	int width = 0, height = 0, spp = 1, maxPathLength = 5;
	int shardIndex = 0, shardCount = 1, sampleOffset = 0, jitter = 1;
	uint32_t pathsPerPass = 32u << 20;
	uint32_t capacity = 0; // rays per buffer
	DeviceBuffer rayO[2], rayD[2], state[2], hits, conO, conD, conE, counters, accumOwn, frame, materials;
	float4* accumExternal = nullptr;
	uint32_t materialCount = 0;
	float lightPos[3] = { 0, 0, 0 }, lightRadiance[3] = { 0, 0, 0 }, sky[3] = { 0, 0, 0 };
	bool hasLight = false;
	uint32_t samplesTaken = 0, frameCounter = 0;
	int probeX = -1, probeY = -1;
	cudaEvent_t ev[2 * (3 * MAX_PATH_LENGTH + 2)] = {};
	float4* accumulator() { return accumExternal ? accumExternal : (float4*)accumOwn.ptr; }
	~WavefrontState()
	{
		DeviceBuffer* all[] = { &rayO[0], &rayO[1], &rayD[0], &rayD[1], &state[0], &state[1], &hits, &conO, &conD, &conE, &counters, &accumOwn, &frame, &materials };
		for (DeviceBuffer* b : all) b->release();
		for (cudaEvent_t e : ev) if (e) cudaEventDestroy( e );
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
	if (!strcmp( name, "pathsPerPass" )) { if (value < 1024) return fail( "Setting pathsPerPass: >= 1024" ); w->pathsPerPass = (uint32_t)value; return 0; }
	return fail( std::string( "Setting: unknown key " ) + name );
}

// CoreAPI_Base::SetTarget (core_api_base.h:90): only width/height/spp are used (lib/rendercore_optix7/rendercore.cpp:286-330)
int RenderCore::SetTarget( int width, int height, int spp )
{
	if (width <= 0 || height <= 0 || spp <= 0) return fail( "SetTarget: width, height and spp must be positive" );
	WavefrontState* w = state( wf );
	w->width = width, w->height = height, w->spp = spp;
	const size_t px = (size_t)width * height;
	cudaError_t e;
	if ((e = w->accumOwn.reserve( px * 16, false, stream )) != cudaSuccess || (e = w->frame.reserve( px * 16, false, stream )) != cudaSuccess) return fail( "SetTarget buffers", e );
	if ((e = w->counters.reserve( sizeof( WaveCounters ) * (MAX_PATH_LENGTH + 2), false, stream )) != cudaSuccess) return fail( "SetTarget counters", e );
	if ((e = cudaMemsetAsync( w->accumulator(), 0, px * 16, stream )) != cudaSuccess) return fail( "SetTarget clear", e );
	w->samplesTaken = 0;
	for (cudaEvent_t& ev : w->ev) if (!ev && (e = cudaEventCreate( &ev )) != cudaSuccess) return fail( "cudaEventCreate", e );
	return 0;
}

int RenderCore::SetAccumulatorBuffer( void* devicePtr )
{
	state( wf )->accumExternal = (float4*)devicePtr;
	return 0;
}

int RenderCore::SetPointLight( const float* pos, const float* radiance )
{
	WavefrontState* w = state( wf );
	if (!pos || !radiance) { w->hasLight = false; return 0; }
	memcpy( w->lightPos, pos, 12 ), memcpy( w->lightRadiance, radiance, 12 );
	w->hasLight = true;
	return 0;
}

int RenderCore::SetSky( const float* radiance )
{
	if (!radiance) return fail( "SetSky: null" );
	memcpy( state( wf )->sky, radiance, 12 );
	return 0;
}

// CoreAPI_Base::SetMaterials (core_api_base.h:100): CoreMaterial[count], 688 bytes each; the diffuse stage keeps color.value
// (offset 0, common_classes.h:177-238)
int RenderCore::SetMaterials( const void* coreMaterials, int count )
{
	WavefrontState* w = state( wf );
	if (count < 0 || (count > 0 && !coreMaterials)) return fail( "SetMaterials: bad arguments" );
	std::vector<float> colors( (size_t)count * 4 );
	for (int i = 0; i < count; i++)
	{
		memcpy( &colors[(size_t)i * 4], (const char*)coreMaterials + (size_t)i * 688, 12 );
		colors[(size_t)i * 4 + 3] = 0;
	}
	return SetMaterialColors( colors.data(), count );
}

int RenderCore::SetMaterialColors( const float* rgba, int count )
{
	WavefrontState* w = state( wf );
	cudaError_t e;
	if ((e = w->materials.reserve( (size_t)(count + 1) * 16, false, stream )) != cudaSuccess) return fail( "materials", e );
	if (count && (e = cudaMemcpyAsync( w->materials.ptr, rgba, (size_t)count * 16, cudaMemcpyHostToDevice, stream )) != cudaSuccess) return fail( "materials upload", e );
	if ((e = cudaStreamSynchronize( stream )) != cudaSuccess) return fail( "materials sync", e );
	w->materialCount = (uint32_t)count;
	return 0;
}

int RenderCore::SetProbe( int x, int y )
{
	state( wf )->probeX = x, state( wf )->probeY = y;
	return 0;
}

// CoreAPI_Base::Render (core_api_base.h:94).  converge: 0 = Converge (keep accumulating), 1 = Restart (clear first)
// (common_classes.h:38-42; lib/rendercore_optix7/rendercore.cpp:627-634).
int RenderCore::Render( const void* viewPyramid, int converge )
{
	WavefrontState* w = wf;
	if (!w || w->width == 0) return fail( "Render: SetTarget has not been called" );
	if (!viewPyramid) return fail( "Render: null view" );
	if (ensureToplevel()) return 1;
	cudaError_t e;
	const float* vp = (const float*)viewPyramid; // pos, p1, p2, p3 (4 x float3), then 5 floats (common_classes.h:362-385)
	FrameParams f;
	memset( &f, 0, sizeof( f ) );
	for (int a = 0; a < 3; a++) f.pos[a] = vp[a], f.p1[a] = vp[3 + a], f.right[a] = vp[6 + a] - vp[3 + a], f.up[a] = vp[9 + a] - vp[3 + a];
	memcpy( f.lightPos, w->lightPos, 12 ), memcpy( f.lightRadiance, w->lightRadiance, 12 ), memcpy( f.sky, w->sky, 12 );
	f.epsilon = epsilon, f.width = (uint32_t)w->width, f.height = (uint32_t)w->height, f.pixelCount = f.width * f.height;
	f.tilesX = (f.width + TILE_W - 1) / TILE_W, f.tilesY = (f.height + TILE_H - 1) / TILE_H;
	const uint32_t nTiles = f.tilesX * f.tilesY;
	f.shardIndex = (uint32_t)w->shardIndex, f.shardCount = (uint32_t)w->shardCount;
	f.ownTiles = nTiles > f.shardIndex ? (nTiles - f.shardIndex + f.shardCount - 1) / f.shardCount : 0;
	f.hasLight = w->hasLight, f.jitter = (uint32_t)w->jitter, f.materialCount = w->materialCount;
	f.probePixel = (w->probeX >= 0 && w->probeY >= 0 && w->probeX < w->width && w->probeY < w->height) ? w->probeX + w->probeY * w->width : -1;
	const size_t px = (size_t)f.pixelCount;
	if (converge == 1 || w->samplesTaken == 0)
	{
		if ((e = cudaMemsetAsync( w->accumulator(), 0, px * 16, stream )) != cudaSuccess) return fail( "accumulator clear", e );
		w->samplesTaken = 0;
	}
	// paths of this shard per sample; several samples share one pass while they fit
	const uint32_t perSample = f.ownTiles * TILE_PIXELS;
	uint32_t samplesPerPass = perSample ? w->pathsPerPass / perSample : 1;
	if (samplesPerPass < 1) samplesPerPass = 1;
	if (samplesPerPass > (uint32_t)w->spp) samplesPerPass = (uint32_t)w->spp;
	const uint32_t cap = perSample * samplesPerPass + 32;
	if (cap > w->capacity)
	{
		DeviceBuffer* bufs[] = { &w->rayO[0], &w->rayO[1], &w->rayD[0], &w->rayD[1], &w->state[0], &w->state[1], &w->hits, &w->conO, &w->conD, &w->conE };
		for (DeviceBuffer* b : bufs) if ((e = b->reserve( (size_t)cap * 16, false, stream )) != cudaSuccess) return fail( "path buffers", e );
		w->capacity = cap;
	}
	WaveCounters* dCounters = (WaveCounters*)w->counters.ptr;
	const int maxLen = w->maxPathLength;
	memset( &stats.totalRays, 0, (char*)&stats.probedInstid - (char*)&stats.totalRays );
	cudaEvent_t frameA = w->ev[0], frameB = w->ev[1];
	cudaEventRecord( frameA, stream );
	for (uint32_t s0 = 0; s0 < (uint32_t)w->spp; s0 += samplesPerPass)
	{
		f.samplesThisPass = (uint32_t)w->spp - s0 < samplesPerPass ? (uint32_t)w->spp - s0 : samplesPerPass;
		// pathIdx = pixel + sample * pixelCount must stay below 2^32: the sample index wraps after `period` samples and the wrap
		// count goes into the salt instead ("samplesTaken * 7907 + pathLength * 91771" in the reference's shade call)
		const uint32_t globalSample = (uint32_t)w->sampleOffset + w->samplesTaken + s0;
		uint32_t period = 0xffffffffu / f.pixelCount - (uint32_t)w->spp;
		if (period < 1 || period > 0x7fffffffu) period = 1;
		f.sampleBase = globalSample % period;
		f.R0 = (globalSample / period + 1) * 7907u;
		if ((e = cudaMemsetAsync( dCounters, 0, sizeof( WaveCounters ) * (MAX_PATH_LENGTH + 2), stream )) != cudaSuccess) return fail( "counter clear", e );
		const uint32_t paths = perSample * f.samplesThisPass;
		if (!paths) continue;
		generateEyeRaysKernel<<<smCount * 4, 256, 0, stream>>>( f, (float4*)w->rayO[0].ptr, (float4*)w->rayD[0].ptr, (float4*)w->state[0].ptr, dCounters );
		stats.kernelLaunches++;
		int in = 0;
		cudaEvent_t* ev = w->ev + 2;
		for (int len = 1; len <= maxLen; len++)
		{
			const int out = in ^ 1;
			cudaEventRecord( ev[0], stream );
			if (TraceClosestDevice( w->rayO[in].ptr, w->rayD[in].ptr, (int)paths, w->hits.ptr, &dCounters[len].extensionRays )) return 1;
			cudaEventRecord( ev[1], stream );
			shadeKernel<<<smCount * 8, 128, 0, stream>>>( f, len, maxLen, (const float4*)w->rayO[in].ptr, (const float4*)w->rayD[in].ptr,
				(const float4*)w->state[in].ptr, (const float4*)w->hits.ptr, (float4*)w->rayO[out].ptr, (float4*)w->rayD[out].ptr, (float4*)w->state[out].ptr,
				(float4*)w->conO.ptr, (float4*)w->conD.ptr, (float4*)w->conE.ptr, (const float4*)shadePool.ptr, (const float4*)instShadePool.ptr,
				(const float4*)w->materials.ptr, w->accumulator(), dCounters );
			stats.kernelLaunches++;
			cudaEventRecord( ev[2], stream );
			if (w->hasLight)
			{
				if (TraceConnectDevice( w->conO.ptr, w->conD.ptr, (int)paths, w->conE.ptr, w->accumulator(), &dCounters[len].shadowRays )) return 1;
			}
			cudaEventRecord( ev[3], stream );
			ev += 4;
			in = out;
		}
		// per-pass statistics (one synchronisation per pass, not per bounce)
		WaveCounters hc[MAX_PATH_LENGTH + 2];
		if ((e = cudaMemcpyAsync( hc, dCounters, sizeof( hc ), cudaMemcpyDeviceToHost, stream )) != cudaSuccess) return fail( "counter read-back", e );
		if ((e = cudaStreamSynchronize( stream )) != cudaSuccess) return fail( "Render", e );
		ev = w->ev + 2;
		for (int len = 1; len <= maxLen; len++, ev += 4)
		{
			float tExt = 0, tShade = 0, tShadow = 0;
			cudaEventElapsedTime( &tExt, ev[0], ev[1] ), cudaEventElapsedTime( &tShade, ev[1], ev[2] ), cudaEventElapsedTime( &tShadow, ev[2], ev[3] );
			const uint32_t n = hc[len].extensionRays;
			if (len == 1) stats.primaryRayCount += n, stats.traceTime0 += tExt;
			else if (len == 2) stats.bounce1RayCount += n, stats.traceTime1 += tExt;
			else stats.deepRayCount += n, stats.traceTimeX += tExt;
			stats.totalExtensionRays += n, stats.totalShadowRays += hc[len].shadowRays;
			stats.shadowTraceTime += tShadow, stats.shadeTime += tShade;
		}
		if (f.probePixel >= 0 && s0 == 0 && hc[0].pad0)
		{
			stats.probedInstid = (int32_t)hc[0].extensionRays, stats.probedTriid = (int32_t)hc[0].shadowRays;
			memcpy( &stats.probedDist, &hc[0].pad0, 4 );
		}
	}
	w->samplesTaken += (uint32_t)w->spp;
	finalizeKernel<<<(f.pixelCount + 255) / 256, 256, 0, stream>>>( w->accumulator(), (float4*)w->frame.ptr, f.pixelCount, 1.0f / (float)w->samplesTaken );
	stats.kernelLaunches++;
	cudaEventRecord( frameB, stream );
	if ((e = cudaStreamSynchronize( stream )) != cudaSuccess) return fail( "Render", e );
	cudaEventElapsedTime( &stats.renderTime, frameA, frameB );
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

void* RenderCore::DeviceAccumulator() { return wf ? (void*)wf->accumulator() : nullptr; }

} // namespace b200
