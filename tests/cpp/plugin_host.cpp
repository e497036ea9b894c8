// plugin_host.cpp -- TEST INFRASTRUCTURE: a stand-in for the reference's RenderSystem that loads RenderCore_B200 the way
// CoreAPI_Base::CreateCoreAPI does (lib/RenderSystem/core_api_base.cpp:97-132: dlopen RTLD_NOW|RTLD_GLOBAL, dlsym CreateCore /
// DestroyCore, CreateCore(), Init()) and then issues the per-frame call order of RenderSystem::SynchronizeSceneData / Render
// (lib/RenderSystem/rendersystem.cpp:214-238) through the C++ vtable only.  The scene comes from a binary file written by the
// pytest side (tests/test_plugin_gpu.py); the frame and the CoreStats go back the same way.
//
//   plugin_host <libRenderCore_B200.so> <scene.bin> <out.bin>
//
// Two builds of this file:
//   * default: the boundary types come from include/lh2_abi.h (the repo's restatement), the library is dlopen'ed by path -- builds anywhere;
//   * -DHOST_USES_REFERENCE_HEADERS (oracle/Makefile -> oracle/_ref/plugin_host_ref, only where /root/reference exists): the boundary types
//     are the REFERENCE'S OWN declarations (lib/RenderSystem/rendersystem.h: CoreAPI_Base, CoreTri, CoreMaterial, ViewPyramid, CoreStats
//     ...) and the core is loaded by the reference's own loader, lib/RenderSystem/core_api_base.cpp compiled in place and linked in:
//     CoreAPI_Base::CreateCoreAPI( "RenderCore_B200" ) (:119-132).  So the vtable order, CoreStats returned by value, the double Init and
//     every struct layout are seen through the reference's compiler view, not through ours.  argv[1] is ignored (the loader dlopens
//     "libRenderCore_B200.so" from LD_LIBRARY_PATH, core_api_base.cpp:97-110).
#ifdef HOST_USES_REFERENCE_HEADERS
#include "rendersystem.h"
#include <cstdarg>
typedef float3 float3_t;
typedef float4 float4_t;
typedef mat4 mat4_t;
typedef int2 int2_t;
typedef uchar4 uchar4_t;
// the one symbol core_api_base.cpp needs from lib/platform/system.cpp (:221-236: message, then exit)
void FatalError( const char* fmt, ... )
{
	va_list args;
	va_start( args, fmt );
	vfprintf( stderr, fmt, args );
	va_end( args );
	exit( 1 );
}
#else
#include "../../include/lh2_abi.h"
using namespace lh2abi;
#endif
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <dlfcn.h>
#include <vector>

static std::vector<char> readAll( const char* path )
{
	FILE* f = fopen( path, "rb" );
	if (!f) { fprintf( stderr, "cannot open %s\n", path ); exit( 2 ); }
	fseek( f, 0, SEEK_END );
	long n = ftell( f );
	fseek( f, 0, SEEK_SET );
	std::vector<char> b( n );
	if (fread( b.data(), 1, n, f ) != (size_t)n) exit( 2 );
	fclose( f );
	return b;
}

struct Reader {
	const char* p;
	template <class T> T get() { T v; memcpy( &v, p, sizeof( T ) ); p += sizeof( T ); return v; }
	const void* bytes( size_t n ) { const void* q = p; p += n; return q; }
};

int main( int argc, char** argv )
{
	if (argc < 4) return 2;
#ifdef HOST_USES_REFERENCE_HEADERS
	CoreAPI_Base* api = CoreAPI_Base::CreateCoreAPI( "RenderCore_B200" ); // the reference's loader: dlopen, dlsym CreateCore, createCore(), Init()
	void* lib = dlopen( "libRenderCore_B200.so", RTLD_NOW | RTLD_NOLOAD ); // the handle the loader keeps to itself, for the read-back symbols below
	if (!api || !lib) { fprintf( stderr, "CreateCoreAPI failed\n" ); return 3; }
#else
	void* lib = dlopen( argv[1], RTLD_NOW | RTLD_GLOBAL );
	if (!lib) { fprintf( stderr, "dlopen: %s\n", dlerror() ); return 3; }
	typedef CoreAPI_Base* (*createCoreFunction)();
	typedef void (*destroyCoreFunction)();
	createCoreFunction createCore = (createCoreFunction)dlsym( lib, "CreateCore" );
	destroyCoreFunction destroyCore = (destroyCoreFunction)dlsym( lib, "DestroyCore" );
	if (!createCore || !destroyCore) { fprintf( stderr, "CreateCore / DestroyCore not exported\n" ); return 3; }
	CoreAPI_Base* api = createCore();
	api->Init(); // the loader calls Init a second time (core_api_base.cpp:128-129)
#endif

	std::vector<char> file = readAll( argv[2] );
	Reader r{ file.data() };
	const int width = r.get<int>(), height = r.get<int>(), spp = r.get<int>(), frames = r.get<int>();
	const int probeX = r.get<int>(), probeY = r.get<int>();
	ViewPyramid view = r.get<ViewPyramid>();
	// SynchronizeSceneData order: sky, textures, materials, geometry, instances, toplevel, lights (rendersystem.cpp:214-222)
	const int skyW = r.get<int>(), skyH = r.get<int>();
	const float3_t* sky = (const float3_t*)r.bytes( (size_t)skyW * skyH * 12 );
	mat4_t identity;
	memset( &identity, 0, sizeof( identity ) );
	float* cells = (float*)&identity;
	cells[0] = cells[5] = cells[10] = cells[15] = 1;
	if (skyW) api->SetSkyData( sky, (uint)skyW, (uint)skyH, identity );
	const int textureCount = r.get<int>();
	std::vector<CoreTexDesc> textures( textureCount );
	for (int i = 0; i < textureCount; i++)
	{
		memset( &textures[i], 0, sizeof( CoreTexDesc ) );
		textures[i].width = (uint)r.get<int>(), textures[i].height = (uint)r.get<int>();
		textures[i].pixelCount = (uint)r.get<int>(), textures[i].MIPlevels = (uint)r.get<int>();
		textures[i].storage = ARGB32;
		textures[i].idata = (uchar4_t*)r.bytes( (size_t)textures[i].pixelCount * 4 );
	}
	api->SetTextures( textures.data(), textureCount ); // before the materials (core_api_base.h:99)
	const int materialCount = r.get<int>();
	std::vector<CoreMaterial> materials( materialCount );
	for (int i = 0; i < materialCount; i++)
	{
		memset( &materials[i], 0, sizeof( CoreMaterial ) );
		// the table of b200_set_material_table: rgb, roughness | transmission, eta, flags, - | extinction rgb, texture | uv scale, uv offset
		const float* c = (const float*)r.bytes( 64 );
		materials[i].color.value.x = c[0], materials[i].color.value.y = c[1], materials[i].color.value.z = c[2];
		// empty map slots carry textureID -1 (common_classes.h:183,190), which is how a HostMaterial arrives
		materials[i].detailColor.textureID = materials[i].normals.textureID = materials[i].detailNormals.textureID = -1;
		materials[i].roughness.textureID = -1;
		materials[i].color.textureID = (int)c[11];
		materials[i].color.uvscale.x = c[12], materials[i].color.uvscale.y = c[13], materials[i].color.uvoffset.x = c[14], materials[i].color.uvoffset.y = c[15];
		materials[i].roughness.value = c[3], materials[i].transmission.value = c[4], materials[i].eta.value = c[5];
		materials[i].flags = (uint)c[6]; // HostMaterial::SMOOTH = 1, HASALPHA = 2
		// the cores keep 1 - absorption (rendercore_optix7/rendercore.cpp:529-531); the test values are dyadic, so this is exact
		materials[i].absorption.value.x = 1 - c[8], materials[i].absorption.value.y = 1 - c[9], materials[i].absorption.value.z = 1 - c[10];
	}
	api->SetMaterials( materials.data(), materialCount );
	const int meshCount = r.get<int>();
	for (int m = 0; m < meshCount; m++)
	{
		const int triCount = r.get<int>();
		const float4_t* verts = (const float4_t*)r.bytes( (size_t)triCount * 48 );
		const CoreTri* tris = (const CoreTri*)r.bytes( (size_t)triCount * sizeof( CoreTri ) );
		api->SetGeometry( m, verts, triCount * 3, triCount, tris, nullptr );
	}
	const int instanceCount = r.get<int>();
	for (int i = 0; i < instanceCount; i++)
	{
		const int mesh = r.get<int>();
		mat4_t T = r.get<mat4_t>();
		api->SetInstance( i, mesh, T );
	}
	api->SetInstance( instanceCount, -1, identity ); // end of the instance stream (rendersystem.cpp:167-169)
	api->UpdateToplevel();
	const int nArea = r.get<int>(), nPoint = r.get<int>(), nSpot = r.get<int>(), nDir = r.get<int>();
	const CoreLightTri* area = (const CoreLightTri*)r.bytes( (size_t)nArea * sizeof( CoreLightTri ) );
	const CorePointLight* point = (const CorePointLight*)r.bytes( (size_t)nPoint * sizeof( CorePointLight ) );
	const CoreSpotLight* spot = (const CoreSpotLight*)r.bytes( (size_t)nSpot * sizeof( CoreSpotLight ) );
	const CoreDirectionalLight* dir = (const CoreDirectionalLight*)r.bytes( (size_t)nDir * sizeof( CoreDirectionalLight ) );
	api->SetLights( area, nArea, point, nPoint, spot, nSpot, dir, nDir );

	// (raw storage: the reference's GLTexture has no default constructor and its real ones call into GL; cores read ID / width / height only)
	alignas( GLTexture ) static char targetStorage[sizeof( GLTexture )];
	GLTexture& target = *reinterpret_cast<GLTexture*>( targetStorage );
	target.ID = 0, target.width = (uint)width, target.height = (uint)height;
	api->SetTarget( &target, (uint)spp );
	int2_t probe;
	probe.x = probeX, probe.y = probeY;
	api->SetProbePos( probe );
	// the six settings RenderSystem::Render sends every frame (rendersystem.cpp:228-236), then B200's own keys
	api->Setting( "epsilon", 1e-4f ), api->Setting( "clampValue", 10.0f ), api->Setting( "clampDirect", 10.0f ), api->Setting( "clampIndirect", 10.0f );
	api->Setting( "filter", 0 ), api->Setting( "TAA", 0 );
	api->Setting( "jitter", 0 ), api->Setting( "maxPathLength", 4 );
	CoreStats stats;
	for (int f = 0; f < frames; f++)
	{
		api->Render( view, f == 0 ? Restart : Converge );
		stats = api->GetCoreStats();
	}
	// pixels: the reference cores write a GL texture; this core keeps the frame in a device buffer with an extern "C" read-back
	typedef void* (*pluginCoreFunction)();
	typedef int (*readFunction)( void*, float* );
	pluginCoreFunction pluginCore = (pluginCoreFunction)dlsym( lib, "b200_plugin_core" );
	readFunction readFrame = (readFunction)dlsym( lib, "b200_read_frame" ), readAccumulator = (readFunction)dlsym( lib, "b200_read_accumulator" );
	if (!pluginCore || !readFrame || !readAccumulator) return 4;
	std::vector<float> frame( (size_t)width * height * 4 ), accumulator( (size_t)width * height * 4 );
	if (readFrame( pluginCore(), frame.data() ) || readAccumulator( pluginCore(), accumulator.data() )) return 5;
	FILE* out = fopen( argv[3], "wb" );
	if (!out) return 6;
	fwrite( frame.data(), 4, frame.size(), out );
	fwrite( accumulator.data(), 4, accumulator.size(), out );
	fwrite( &stats, sizeof( stats ), 1, out );
	const char* name = stats.deviceName ? stats.deviceName : "";
	char nameBuf[64] = {};
	strncpy( nameBuf, name, 63 );
	fwrite( nameBuf, 1, 64, out );
	fclose( out );
	api->Shutdown();
#ifndef HOST_USES_REFERENCE_HEADERS
	destroyCore(); // (the reference looks DestroyCore up and never calls it, core_api_base.cpp:126)
#endif
	return 0;
}
