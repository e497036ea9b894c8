// core_api.cpp -- the two faces of libRenderCore_B200.so:
//   (1) class CoreAPI : CoreAPI_Base + extern "C" CreateCore/DestroyCore -- what RenderSystem loads with
//       dlopen/dlsym (lib/RenderSystem/core_api_base.cpp:97-132); same shape as the reference cores'
//       core_api.cpp (lib/RenderCore_Bart/core_api.cpp:18-101): a forwarder to one RenderCore singleton.
//   (2) the flat b200_* C ABI declared in include/rendercore_b200.h.
// Errors: the C ABI returns non-zero + b200_last_error(); the C++ face keeps the reference's convention
// (void methods, fatal on error: lib/platform/system.cpp:221-236 prints and exits).
#include "rendercore.h"
#include "../../include/lh2_abi.h"
#include <cstdio>
#include <cstdlib>

using namespace b200;

struct b200_core { RenderCore* rc; };

static thread_local std::string gLastError;

#define GUARD( c ) if (!(c) || !(c)->rc) { gLastError = "null core handle"; return 1; }
#define FORWARD( c, call ) GUARD( c ) { int rcode = (c)->rc->call; if (rcode) gLastError = (c)->rc->Error(); return rcode; }

extern "C" {

int b200_create( b200_core** out, int device )
{
	if (!out) { gLastError = "b200_create: null out pointer"; return 1; }
	*out = nullptr;
	RenderCore* rc = new RenderCore( device );
	if (!rc->ok()) { gLastError = rc->Error(); delete rc; return 1; }
	*out = new b200_core{ rc };
	return 0;
}
void b200_destroy( b200_core* c ) { if (c) { delete c->rc; delete c; } }
const char* b200_last_error( void ) { return gLastError.c_str(); }
int b200_set_stream( b200_core* c, void* stream ) { FORWARD( c, SetStream( stream ) ) }

int b200_set_geometry( b200_core* c, int meshIdx, const float* vertexData, int vertexCount, int triangleCount, const void* triangles,
	const uint32_t* alphaFlags ) { FORWARD( c, SetGeometry( meshIdx, vertexData, vertexCount, triangleCount, triangles, alphaFlags ) ) }
int b200_set_instance( b200_core* c, int instanceIdx, int modelIdx, const float* transform ) { FORWARD( c, SetInstance( instanceIdx, modelIdx, transform ) ) }
int b200_update_toplevel( b200_core* c ) { FORWARD( c, UpdateToplevel() ) }
int b200_refit( b200_core* c, int meshIdx, const float* vertexData, int vertexCount ) { FORWARD( c, Refit( meshIdx, vertexData, vertexCount ) ) }
int b200_refit_device( b200_core* c, int meshIdx, const void* dVerts, int vertexCount ) { FORWARD( c, RefitDevice( meshIdx, dVerts, vertexCount, false ) ) }
int b200_setting( b200_core* c, const char* name, float value ) { FORWARD( c, Setting( name ? name : "", value ) ) }

int b200_trace_closest( b200_core* c, const float* O4, const float* D4, int n, float* hits ) { FORWARD( c, TraceClosest( O4, D4, n, hits ) ) }
int b200_trace_any( b200_core* c, const float* O4, const float* D4, int n, uint32_t* bits ) { FORWARD( c, TraceAny( O4, D4, n, bits ) ) }
int b200_trace_closest_async( b200_core* c, const float* O4, const float* D4, int n, float* hits ) { FORWARD( c, TraceClosestAsync( O4, D4, n, hits ) ) }
int b200_trace_any_async( b200_core* c, const float* O4, const float* D4, int n, uint32_t* bits ) { FORWARD( c, TraceAnyAsync( O4, D4, n, bits ) ) }
int b200_wait( b200_core* c ) { FORWARD( c, WaitWaves() ) }
int b200_trace_closest_noalloc( b200_core* c, const float* O4, const float* D4, int n, float* hits ) { FORWARD( c, TraceClosestNoAlloc( O4, D4, n, hits ) ) }
int b200_trace_closest_device( b200_core* c, const void* dO4, const void* dD4, int n, void* dHits ) { FORWARD( c, TraceClosestDevice( dO4, dD4, n, dHits ) ) }
int b200_trace_any_device( b200_core* c, const void* dO4, const void* dD4, int n, void* dBits ) { FORWARD( c, TraceAnyDevice( dO4, dD4, n, dBits ) ) }
int b200_read_step_counters( b200_core* c, int n, uint32_t* nodes, uint32_t* tris ) { FORWARD( c, ReadStepCounters( n, nodes, tris ) ) }
int b200_read_tree( b200_core* c, int which, void* dst, size_t bytes ) { FORWARD( c, ReadTree( which, dst, bytes ) ) }

// ---- the filter's stages on caller-owned DEVICE buffers (filter.cu); current device, default stream, synchronous
static int filterDone( const char* what, cudaError_t e )
{
	if (e == cudaSuccess) e = cudaDeviceSynchronize();
	if (e != cudaSuccess) { gLastError = std::string( what ) + ": " + cudaGetErrorString( e ); return 1; }
	return 0;
}
static FilterFrame filterFrame( const void* accumulator, void* features, const void* worldPos, const void* prevWorldPos, const void* deltaDepth, void* shading,
	void* motion, void* moments, const void* prevMoments, int w, int h )
{
	FilterFrame f;
	f.accumulator = (const float4*)accumulator, f.features = (uint4*)features, f.worldPos = (const float4*)worldPos, f.prevWorldPos = (const float4*)prevWorldPos;
	f.deltaDepth = (const float4*)deltaDepth, f.shading = (float4*)shading, f.motion = (float2*)motion, f.moments = (float4*)moments;
	f.prevMoments = (const float4*)prevMoments, f.width = w, f.height = h;
	return f;
}
int b200_filter_prepare( const void* accumulator, void* features, const void* worldPos, const void* prevWorldPos, void* shading, void* motion, void* moments,
	const void* prevMoments, const void* deltaDepth, const float* prevView, float j0, float j1, float prevj0, float prevj1, int w, int h, unsigned spp,
	float directClamp, float indirectClamp, int camIsStationary )
{
	if (!accumulator || !features || !worldPos || !prevWorldPos || !shading || !motion || !moments || !prevMoments || !deltaDepth || !prevView || w < 1 || h < 1 || !spp)
	{ gLastError = "b200_filter_prepare: bad arguments"; return 1; }
	const FilterFrame f = filterFrame( accumulator, features, worldPos, prevWorldPos, deltaDepth, shading, motion, moments, prevMoments, w, h );
	return filterDone( "b200_filter_prepare", LaunchFilterPrepare( f, MakeFilterReprojection( prevView, h, j0, j1, prevj0, prevj1, spp, directClamp, indirectClamp, camIsStationary ), 0 ) );
}
int b200_filter_reconstruct( const void* features, const void* prevWorldPos, const void* worldPos, const void* deltaDepth, const void* motion, const void* moments,
	const void* A, const void* B, void* C, int w, int h, int phase, int lastPass )
{
	if (!features || !prevWorldPos || !worldPos || !deltaDepth || !motion || !moments || !A || !C || (phase == 1 && !B) || w < 1 || h < 1 || phase < 1 || phase > 8)
	{ gLastError = "b200_filter_reconstruct: bad arguments"; return 1; }
	const FilterFrame f = filterFrame( nullptr, (void*)features, worldPos, prevWorldPos, deltaDepth, nullptr, (void*)motion, (void*)moments, nullptr, w, h );
	return filterDone( "b200_filter_reconstruct", LaunchFilterReconstruct( f, (const float4*)A, (const float4*)B, (float4*)C, phase, lastPass, 0 ) );
}
int b200_filter_taa( const void* pixels, const void* prevPixels, const void* motion, void* out, int w, int h )
{
	if (!pixels || !prevPixels || !motion || !out || out == pixels || w < 1 || h < 1) { gLastError = "b200_filter_taa: bad arguments (out must not be pixels)"; return 1; }
	return filterDone( "b200_filter_taa", LaunchFilterTaa( (const float4*)pixels, (const float4*)prevPixels, (const float2*)motion, (float4*)out, w, h, 0 ) );
}
int b200_filter_unsharpen( const void* pixels, void* frame, int w, int h )
{
	if (!pixels || !frame || w < 1 || h < 1) { gLastError = "b200_filter_unsharpen: bad arguments"; return 1; }
	return filterDone( "b200_filter_unsharpen", LaunchFilterUnsharpen( (const float4*)pixels, (float4*)frame, w, h, 0 ) );
}
int b200_filter_finish_no_taa( const void* pixels, void* frame, int w, int h )
{
	if (!pixels || !frame || w < 1 || h < 1) { gLastError = "b200_filter_finish_no_taa: bad arguments"; return 1; }
	return filterDone( "b200_filter_finish_no_taa", LaunchFilterFinishNoTaa( (const float4*)pixels, (float4*)frame, w, h, 0 ) );
}

int b200_set_target( b200_core* c, int w, int h, int spp ) { FORWARD( c, SetTarget( w, h, spp ) ) }
int b200_set_point_light( b200_core* c, const float* p, const float* r ) { FORWARD( c, SetPointLight( p, r ) ) }
int b200_set_sky( b200_core* c, const float* r ) { FORWARD( c, SetSky( r ) ) }
int b200_set_lights( b200_core* c, const void* area, int nArea, const void* point, int nPoint, const void* spot, int nSpot, const void* dir, int nDir )
{
	FORWARD( c, SetLights( area, nArea, point, nPoint, spot, nSpot, dir, nDir ) )
}
int b200_set_sky_data( b200_core* c, const float* rgb, int width, int height ) { FORWARD( c, SetSkyData( rgb, width, height ) ) }
int b200_set_materials( b200_core* c, const void* coreMaterials, int count ) { FORWARD( c, SetMaterials( coreMaterials, count ) ) }
int b200_set_material_colors( b200_core* c, const float* rgba, int count ) { FORWARD( c, SetMaterialColors( rgba, count ) ) }
int b200_set_material_table( b200_core* c, const float* props16, int count ) { FORWARD( c, SetMaterialTable( props16, count ) ) }
int b200_set_textures( b200_core* c, const void* coreTexDescs, int count ) { FORWARD( c, SetTextures( coreTexDescs, count ) ) }
int b200_set_blue_noise( b200_core* c, const uint32_t* words, int count ) { FORWARD( c, SetBlueNoise( words, count ) ) }
int b200_disney_probe( b200_core* c, int mode, int n, const float* in28, float* out8 ) { FORWARD( c, DisneyProbe( mode, n, in28, out8 ) ) }
int b200_set_accumulator_buffer( b200_core* c, void* dAcc ) { FORWARD( c, SetAccumulatorBuffer( dAcc ) ) }
int b200_peer_handle( b200_core* c, void* out128 ) { FORWARD( c, PeerHandle( out128 ) ) }
int b200_open_peers( b200_core* c, const void* handles, int world, int rank ) { FORWARD( c, OpenPeers( handles, world, rank ) ) }
int b200_close_peers( b200_core* c ) { FORWARD( c, ClosePeers() ) }
int b200_gather_finalize( b200_core* c, uint32_t totalSamples ) { FORWARD( c, GatherFinalize( totalSamples ) ) }
int b200_finalize( b200_core* c, uint32_t totalSamples ) { FORWARD( c, Finalize( totalSamples ) ) }
int b200_trace_connect_device( b200_core* c, const void* dO4, const void* dD4, int n, const void* dE4, void* dAcc ) { FORWARD( c, TraceConnectDevice( dO4, dD4, n, dE4, dAcc ) ) }
int b200_render( b200_core* c, const void* view, int converge ) { FORWARD( c, Render( view, converge ) ) }
int b200_read_accumulator( b200_core* c, float* dst ) { FORWARD( c, ReadAccumulator( dst ) ) }
int b200_read_frame( b200_core* c, float* dst ) { FORWARD( c, ReadFrame( dst ) ) }
int b200_read_frame_async( b200_core* c, float* dst ) { FORWARD( c, ReadFrameAsync( dst ) ) }
int b200_wait_frame( b200_core* c ) { FORWARD( c, WaitFrame() ) }
int b200_read_filter_buffer( b200_core* c, int which, void* dst ) { FORWARD( c, ReadFilterBuffer( which, dst ) ) }
void* b200_device_accumulator( b200_core* c ) { return (c && c->rc) ? c->rc->DeviceAccumulator() : nullptr; }
int b200_get_stats( b200_core* c, b200_stats* out ) { GUARD( c ) if (!out) { gLastError = "null stats"; return 1; } c->rc->GetStats( out ); return 0; }
int b200_set_probe( b200_core* c, int x, int y ) { FORWARD( c, SetProbe( x, y ) ) }

void* b200_host_alloc( uint64_t bytes )
{
	void* p = nullptr;
	if (cudaHostAlloc( &p, (size_t)bytes, cudaHostAllocDefault ) != cudaSuccess) { cudaGetLastError(); return nullptr; }
	return p;
}
void b200_host_free( void* p ) { if (p) cudaFreeHost( p ); }

} // extern "C"

// ======================================================================================================
// C++ plugin face
// ======================================================================================================
namespace {

using namespace lh2abi;

RenderCore* gCore = nullptr; // file-static singleton like `static RenderCore* core` (lib/RenderCore_Bart/core_api.cpp:37)
char gDeviceName[256] = "";

[[noreturn]] void Fatal( const char* where, const std::string& msg )
{
	// FatalError convention of the reference (lib/platform/system.cpp:221-236): message, then exit
	fprintf( stderr, "RenderCore_B200 fatal error in %s: %s\n", where, msg.c_str() );
	exit( 1 );
}
#define CHK( call ) do { if (gCore->call) Fatal( #call, gCore->Error() ); } while (0)

class CoreAPI : public CoreAPI_Base {
public:
	CoreStats GetCoreStats() override
	{
		CoreStats s;
		b200_stats b;
		gCore->GetStats( &b );
		s.deviceName = gDeviceName;
		s.SMcount = b.smCount, s.ccMajor = b.ccMajor, s.ccMinor = b.ccMinor, s.VRAM = b.vramMB;
		s.bvhBuildTime = b.bvhBuildTime * 0.001f; // seconds, as rendercore_optix7 reports it
		s.totalRays = b.totalRays, s.totalExtensionRays = b.totalExtensionRays, s.totalShadowRays = b.totalShadowRays;
		s.renderTime = b.renderTime * 0.001f;
		s.primaryRayCount = b.primaryRayCount, s.traceTime0 = b.traceTime0 * 0.001f;
		s.bounce1RayCount = b.bounce1RayCount, s.traceTime1 = b.traceTime1 * 0.001f;
		s.deepRayCount = b.deepRayCount, s.traceTimeX = b.traceTimeX * 0.001f;
		s.shadowTraceTime = b.shadowTraceTime * 0.001f, s.shadeTime = b.shadeTime * 0.001f;
		s.probedInstid = b.probedInstid, s.probedTriid = b.probedTriid, s.probedDist = b.probedDist;
		s.filterTime = b.filterTime * 0.001f;
		s.argb32TexelCount = b.argb32TexelCount, s.argb128TexelCount = b.argb128TexelCount, s.nrm32TexelCount = b.nrm32TexelCount;
		return s;
	}
	void Init() override
	{
		// called twice by the loader (CreateCore, then CreateCoreAPI: core_api_base.cpp:128-129): idempotent
		if (gCore) return;
		int device = 0;
		if (const char* env = getenv( "LOCAL_RANK" )) device = atoi( env ); // one process per GPU under torchrun
		if (const char* env = getenv( "B200_DEVICE" )) device = atoi( env );
		gCore = new RenderCore( device );
		if (!gCore->ok()) Fatal( "Init", gCore->Error() );
		cudaDeviceProp prop;
		if (cudaGetDeviceProperties( &prop, device ) == cudaSuccess) snprintf( gDeviceName, sizeof( gDeviceName ), "%s", prop.name );
	}
	void SetProbePos( const int2_t pos ) override { CHK( SetProbe( pos.x, pos.y ) ); }
	void SetTarget( GLTexture* target, const uint spp ) override
	{
		// headless: only width/height are read (no GL object exists on this box)
		CHK( SetTarget( (int)target->width, (int)target->height, (int)spp ) );
	}
	void Setting( const char* name, float value ) override
	{
		// unknown keys are ignored, as the reference cores ignore them (rendercore_optix7/rendercore.cpp:597-615); a KNOWN key with a value it cannot take
		// is reported -- the method is void and the RenderSystem sends its keys every frame, so: on stderr, the first time each
		if (gCore->Setting( name, value ) && gCore->Error().rfind( "Setting: unknown key", 0 ) != 0)
		{
			static std::string reported;
			const std::string key = std::string( "|" ) + (name ? name : "") + "|";
			if (reported.find( key ) == std::string::npos) reported += key, fprintf( stderr, "RenderCore_B200: %s (value %g ignored)\n", gCore->Error().c_str(), value );
		}
	}
	void Render( const ViewPyramid& view, const Convergence converge ) override { CHK( Render( &view, (int)converge ) ); }
	void Shutdown() override { delete gCore; gCore = nullptr; }
	void SetTextures( const CoreTexDesc* tex, const int textureCount ) override { CHK( SetTextures( tex, textureCount ) ); }
	void SetMaterials( CoreMaterial* mat, const int materialCount ) override { CHK( SetMaterials( mat, materialCount ) ); }
	void SetLights( const CoreLightTri* areaLights, const int areaLightCount, const CorePointLight* pointLights, const int pointLightCount,
		const CoreSpotLight* spotLights, const int spotLightCount, const CoreDirectionalLight* directionalLights, const int directionalLightCount ) override
	{
		CHK( SetLights( areaLights, areaLightCount, pointLights, pointLightCount, spotLights, spotLightCount, directionalLights, directionalLightCount ) );
	}
	void SetSkyData( const float3_t* pixels, const uint width, const uint height, const mat4_t& ) override
	{
		if (!pixels || !width || !height) return;
		CHK( SetSkyData( &pixels[0].x, (int)width, (int)height ) );
	}
	void SetGeometry( const int meshIdx, const float4_t* vertexData, const int vertexCount, const int triangleCount, const CoreTri* triangles,
		const uint* alphaFlags ) override
	{
		CHK( SetGeometry( meshIdx, (const float*)vertexData, vertexCount, triangleCount, triangles, alphaFlags ) );
	}
	void SetInstance( const int instanceIdx, const int modelIdx, const mat4_t& transform ) override
	{
		CHK( SetInstance( instanceIdx, modelIdx, transform.cell ) );
	}
	void UpdateToplevel() override { CHK( UpdateToplevel() ); }
};

CoreAPI* gApi = nullptr;

} // namespace

extern "C" B200_API lh2abi::CoreAPI_Base* CreateCore()
{
	// no gladLoadGL here (lib/RenderCore_Bart/core_api.cpp:23): this core is headless
	if (!gApi)
	{
		gApi = new CoreAPI();
		gApi->Init();
	}
	return gApi;
}

// handle onto the plugin's singleton so headless hosts can read the frame back through the b200_* calls
extern "C" B200_API b200_core* b200_plugin_core( void )
{
	static b200_core handle{ nullptr };
	handle.rc = gCore;
	return gCore ? &handle : nullptr;
}

extern "C" B200_API void DestroyCore()
{
	if (gApi) { gApi->Shutdown(); delete gApi; gApi = nullptr; }
}
