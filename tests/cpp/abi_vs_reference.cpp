// tests/cpp/abi_vs_reference.cpp -- compiled only where /root/reference exists: proves that the layouts
// restated in include/lh2_abi.h equal the reference's own declarations (sizes and the offsets the core reads).
#include "core_settings.h" // lib/RenderCore_Bart: pulls platform.h, core_api_base.h, common_classes.h
#include "lh2_abi.h"
#include <cstddef>
#include <cstdio>
#define SAME_SIZE( T ) static_assert( sizeof( ::T ) == sizeof( lh2abi::T ), #T )
SAME_SIZE( CoreTri );
SAME_SIZE( CoreMaterial );
SAME_SIZE( CoreTexDesc );
SAME_SIZE( CoreLightTri );
SAME_SIZE( CorePointLight );
SAME_SIZE( CoreSpotLight );
SAME_SIZE( CoreDirectionalLight );
SAME_SIZE( ViewPyramid );
SAME_SIZE( CoreStats );
SAME_SIZE( GLTexture );
static_assert( sizeof( mat4 ) == sizeof( lh2abi::mat4_t ), "mat4" );
static_assert( sizeof( ::int2 ) == sizeof( lh2abi::int2_t ) && alignof( ::int2 ) == alignof( lh2abi::int2_t ), "int2" );
static_assert( offsetof( ::CoreTri, material ) == offsetof( lh2abi::CoreTri, material ), "material" );
static_assert( offsetof( ::CoreTri, vertex0 ) == offsetof( lh2abi::CoreTri, vertex0 ), "vertex0" );
static_assert( offsetof( ::CoreTri, vertex2 ) == offsetof( lh2abi::CoreTri, vertex2 ), "vertex2" );
static_assert( offsetof( ::CoreStats, probedInstid ) == offsetof( lh2abi::CoreStats, probedInstid ), "stats" );
static_assert( offsetof( ::CoreStats, traceTimeX ) == offsetof( lh2abi::CoreStats, traceTimeX ), "stats" );
static_assert( offsetof( ::GLTexture, width ) == offsetof( lh2abi::GLTexture, width ), "GLTexture" );
static_assert( offsetof( ::CoreMaterial, reflection ) == offsetof( lh2abi::CoreMaterial, reflection ), "material" );
// every CoreMaterial field the core reads, incl. the padded float2 members of the value structs (common_types.h:62: float2 is ALIGN( 8 ))
#define SAME_OFF( T, F ) static_assert( offsetof( ::T, F ) == offsetof( lh2abi::T, F ), #T "::" #F )
SAME_OFF( CoreMaterial::Vec3Value, textureID ); SAME_OFF( CoreMaterial::Vec3Value, scale ); SAME_OFF( CoreMaterial::Vec3Value, uvscale ); SAME_OFF( CoreMaterial::Vec3Value, uvoffset );
SAME_OFF( CoreMaterial::ScalarValue, textureID ); SAME_OFF( CoreMaterial::ScalarValue, scale ); SAME_OFF( CoreMaterial::ScalarValue, uvscale ); SAME_OFF( CoreMaterial::ScalarValue, uvoffset );
SAME_OFF( CoreMaterial, color ); SAME_OFF( CoreMaterial, detailColor ); SAME_OFF( CoreMaterial, normals ); SAME_OFF( CoreMaterial, detailNormals ); SAME_OFF( CoreMaterial, flags );
SAME_OFF( CoreMaterial, absorption ); SAME_OFF( CoreMaterial, metallic ); SAME_OFF( CoreMaterial, subsurface ); SAME_OFF( CoreMaterial, specular ); SAME_OFF( CoreMaterial, roughness );
SAME_OFF( CoreMaterial, specularTint ); SAME_OFF( CoreMaterial, anisotropic ); SAME_OFF( CoreMaterial, sheen ); SAME_OFF( CoreMaterial, sheenTint ); SAME_OFF( CoreMaterial, clearcoat );
SAME_OFF( CoreMaterial, clearcoatGloss ); SAME_OFF( CoreMaterial, transmission ); SAME_OFF( CoreMaterial, eta );
SAME_OFF( CoreTri, ltriIdx ); SAME_OFF( CoreTri, vN0 ); SAME_OFF( CoreTri, T ); SAME_OFF( CoreTri, B ); SAME_OFF( CoreTri, alpha ); SAME_OFF( CoreTri, LOD ); SAME_OFF( CoreTri, Nx );
SAME_OFF( CoreTexDesc, width ); SAME_OFF( CoreTexDesc, pixelCount ); SAME_OFF( CoreTexDesc, MIPlevels ); SAME_OFF( CoreTexDesc, storage );
SAME_OFF( CoreLightTri, triIdx ); SAME_OFF( CoreLightTri, instIdx ); SAME_OFF( CoreLightTri, radiance ); SAME_OFF( CoreLightTri, area );
static_assert( offsetof( ::ViewPyramid, p3 ) == offsetof( lh2abi::ViewPyramid, p3 ), "view" );
int main() { printf( "abi ok\n" ); return 0; }
