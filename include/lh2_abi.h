// include/lh2_abi.h -- the Lighthouse 2 core boundary, restated for RenderCore_B200.
//
// Every type that crosses the plugin boundary (lib/RenderSystem/core_api_base.h:78-114) is declared
// here with the SAME memory layout as the reference declaration it cites, and the layout is pinned by
// static_asserts (sizes/offsets measured against the reference headers with g++ 13.3, SURVEY.md
// Appendix B).  tests/cpp/abi_check.cpp additionally compiles these asserts against the reference's
// own headers when /root/reference is present.
//
// Only data layout and the virtual-function ORDER are part of the ABI (Itanium C++ ABI: the vtable
// slot of a virtual is its declaration index; CoreAPI_Base has no virtual destructor).
#pragma once
#include <cstddef>
#include <cstdint>
#include <cstring>

namespace lh2abi {

typedef unsigned int uint;

// lib/RenderSystem/common_types.h:55-80 (host-side vector types; plain aggregates)
struct int2_t { int x, y; } __attribute__((aligned(8)));
struct float2_t { float x, y; } __attribute__((aligned(8)));
struct float3_t { float x, y, z; };
struct float4_t { float x, y, z, w; } __attribute__((aligned(16)));
struct uchar4_t { unsigned char x, y, z, w; };

// lib/RenderSystem/common_types.h:469-632 -- row-major cell[16], translation in cells 3,7,11
struct mat4_t {
	float cell[16];
	static mat4_t Identity() { mat4_t m; memset( m.cell, 0, 64 ); m.cell[0] = m.cell[5] = m.cell[10] = m.cell[15] = 1; return m; }
};

// lib/RenderSystem/common_classes.h:38-42
enum Convergence { Converge = 0, Restart = 1 };

// lib/RenderSystem/common_classes.h:57-91 -- 176 bytes; the traversal path reads vertex0/1/2 only
struct CoreTri {
	float u0, u1, u2; int ltriIdx;
	float v0, v1, v2; uint material;
	float3_t vN0; float Nx;
	float3_t vN1; float Ny;
	float3_t vN2; float Nz;
	float3_t T; float area;
	float3_t B; float invArea;
	float3_t alpha; float LOD;
	float3_t vertex0; float dummy0;
	float3_t vertex1; float dummy1;
	float3_t vertex2; float dummy2;
};
static_assert( sizeof( CoreTri ) == 176, "CoreTri layout (common_classes.h:82)" );
static_assert( offsetof( CoreTri, ltriIdx ) == 12 && offsetof( CoreTri, material ) == 28, "CoreTri" );
static_assert( offsetof( CoreTri, Nx ) == 44 && offsetof( CoreTri, Ny ) == 60 && offsetof( CoreTri, Nz ) == 76, "CoreTri" );
static_assert( offsetof( CoreTri, vertex0 ) == 128 && offsetof( CoreTri, vertex1 ) == 144 && offsetof( CoreTri, vertex2 ) == 160, "CoreTri" );

// lib/RenderSystem/common_classes.h:177-238 -- 688 bytes
struct CoreMaterial {
	struct Vec3Value { float3_t value; int textureID; float scale; float2_t uvscale, uvoffset; };
	struct ScalarValue { float value; int textureID; int component; float scale; float2_t uvscale, uvoffset; };
	Vec3Value color, detailColor, normals, detailNormals;
	uint flags;
	Vec3Value absorption;
	ScalarValue metallic, subsurface, specular, roughness, specularTint, anisotropic, sheen, sheenTint, clearcoat, clearcoatGloss,
		transmission, eta;
	ScalarValue reflection, refraction, ior;
};
static_assert( sizeof( CoreMaterial::Vec3Value ) == 40 && sizeof( CoreMaterial::ScalarValue ) == 32, "CoreMaterial values" );
static_assert( sizeof( CoreMaterial ) == 688, "CoreMaterial layout" );
static_assert( offsetof( CoreMaterial, flags ) == 160 && offsetof( CoreMaterial, absorption ) == 168, "CoreMaterial" );
static_assert( offsetof( CoreMaterial, reflection ) == 592 && offsetof( CoreMaterial, ior ) == 656, "CoreMaterial" );
// float2 is ALIGN( 8 ) in the reference (common_types.h:62): four bytes of padding sit in front of Vec3Value::uvscale
static_assert( offsetof( CoreMaterial::Vec3Value, textureID ) == 12 && offsetof( CoreMaterial::Vec3Value, scale ) == 16, "Vec3Value" );
static_assert( offsetof( CoreMaterial::Vec3Value, uvscale ) == 24 && offsetof( CoreMaterial::Vec3Value, uvoffset ) == 32, "Vec3Value" );
static_assert( offsetof( CoreMaterial::ScalarValue, textureID ) == 4 && offsetof( CoreMaterial::ScalarValue, scale ) == 12, "ScalarValue" );
static_assert( offsetof( CoreMaterial::ScalarValue, uvscale ) == 16 && offsetof( CoreMaterial::ScalarValue, uvoffset ) == 24, "ScalarValue" );
static_assert( offsetof( CoreMaterial, detailColor ) == 40 && offsetof( CoreMaterial, normals ) == 80 && offsetof( CoreMaterial, detailNormals ) == 120, "CoreMaterial" );
static_assert( offsetof( CoreMaterial, roughness ) == 304 && offsetof( CoreMaterial, transmission ) == 528 && offsetof( CoreMaterial, eta ) == 560, "CoreMaterial" );

// lib/RenderSystem/common_classes.h:240-269 (host layout)
enum TexelStorage { ARGB32 = 0, ARGB128, NRM32 };
struct CoreTexDesc {
	union { float4_t* fdata; uchar4_t* idata; };
	uint width, height, flags, pixelCount, firstPixel, MIPlevels;
	TexelStorage storage;
};
static_assert( sizeof( CoreTexDesc ) == 40, "CoreTexDesc layout" );

// lib/RenderSystem/common_classes.h:275-360
struct CoreLightTri {
	float3_t centre; float energy;
	float3_t N; float area;
	float3_t radiance; int dummy2;
	float3_t vertex0; int triIdx;
	float3_t vertex1; int instIdx;
	float3_t vertex2; int dummy1;
};
struct CorePointLight { float3_t position; float energy; float3_t radiance; int dummy; };
struct CoreSpotLight { float3_t position; float cosInner; float3_t radiance; float cosOuter; float3_t direction; int dummy; };
struct CoreDirectionalLight { float3_t direction; float energy; float3_t radiance; int dummy; };
static_assert( sizeof( CoreLightTri ) == 96 && sizeof( CorePointLight ) == 32, "light layouts" );
static_assert( sizeof( CoreSpotLight ) == 48 && sizeof( CoreDirectionalLight ) == 32, "light layouts" );

// lib/RenderSystem/common_classes.h:362-385 -- 68 bytes
struct ViewPyramid {
	float3_t pos, p1, p2, p3;
	float aperture, spreadAngle, imagePlane, focalDistance, distortion;
};
static_assert( sizeof( ViewPyramid ) == 68, "ViewPyramid layout" );

// lib/RenderSystem/core_api_base.h:30-61 -- 104 bytes, returned BY VALUE from GetCoreStats
struct CoreStats {
	char* deviceName = 0;
	uint SMcount = 0, ccMajor = 0, ccMinor = 0, VRAM = 0;
	uint argb32TexelCount = 0, argb128TexelCount = 0, nrm32TexelCount = 0;
	float bvhBuildTime = 0;
	uint totalRays = 0, totalExtensionRays = 0, totalShadowRays = 0;
	float renderTime = 0;
	uint primaryRayCount = 0; float traceTime0 = 0;
	uint bounce1RayCount = 0; float traceTime1 = 0;
	uint deepRayCount = 0; float traceTimeX = 0;
	float shadowTraceTime = 0, shadeTime = 0, filterTime = 0;
	int probedInstid = -1, probedTriid = -1; float probedDist = 0;
};
static_assert( sizeof( CoreStats ) == 104, "CoreStats layout" );
static_assert( offsetof( CoreStats, totalRays ) == 40 && offsetof( CoreStats, renderTime ) == 52, "CoreStats" );
static_assert( offsetof( CoreStats, probedInstid ) == 92, "CoreStats" );

// lib/platform/system.h:233-253 -- the cores read only ID/width/height; non-virtual, 12 bytes.
// Headless stand-in: no GL object exists, ID is ignored by RenderCore_B200.
struct GLTexture { uint ID = 0, width = 0, height = 0; };
static_assert( sizeof( GLTexture ) == 12, "GLTexture stand-in" );

// lib/RenderSystem/core_api_base.h:78-114 -- declaration order IS the vtable order.
class CoreAPI_Base {
public:
	virtual CoreStats GetCoreStats() = 0;                                                   // slot 0
	virtual void Init() = 0;                                                                // 1
	virtual void SetProbePos( const int2_t pos ) = 0;                                       // 2
	virtual void SetTarget( GLTexture* target, const uint spp ) = 0;                        // 3
	virtual void Setting( const char* name, float value ) = 0;                              // 4
	virtual void Render( const ViewPyramid& view, const Convergence converge ) = 0;         // 5
	virtual void Shutdown() = 0;                                                            // 6
	virtual void SetTextures( const CoreTexDesc* tex, const int textureCount ) = 0;         // 7
	virtual void SetMaterials( CoreMaterial* mat, const int materialCount ) = 0;            // 8
	virtual void SetLights( const CoreLightTri* areaLights, const int areaLightCount, const CorePointLight* pointLights,
		const int pointLightCount, const CoreSpotLight* spotLights, const int spotLightCount,
		const CoreDirectionalLight* directionalLights, const int directionalLightCount ) = 0; // 9
	virtual void SetSkyData( const float3_t* pixels, const uint width, const uint height, const mat4_t& worldToLight ) = 0; // 10
	virtual void SetGeometry( const int meshIdx, const float4_t* vertexData, const int vertexCount, const int triangleCount,
		const CoreTri* triangles, const uint* alphaFlags = 0 ) = 0;                         // 11
	virtual void SetInstance( const int instanceIdx, const int modelIdx, const mat4_t& transform ) = 0; // 12
	virtual void UpdateToplevel() = 0;                                                      // 13
};

} // namespace lh2abi
