// oracle/ref_shade_driver.cpp -- TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// C-ABI array driver around the UNMODIFIED Lambert BSDF and light-sampling helpers of the reference's GPU cores, compiled for the
// host in place (never copied) from /root/reference by oracle/Makefile into oracle/_ref/liboracle_shade_ref.so:
//   SampleBSDF / EvaluateBSDF, Fr_L, Refract_L                      lib/sharedBSDFs/lambert.h:32-126
//   Potential*LightContribution, CalculateLightPDF, LightPickProb,
//   RandomBarycentrics, RandomPointOnLight                          lib/CUDA/shared_kernel_code/lights_shared.h:33-262
//   ShadingData + parameter macros, Tangent2World, DiffuseReflectionCosWeighted, SafeOrigin ...
//                                                                   lib/CUDA/shared_kernel_code/tools_shared.h
// It pins the checker of SURVEY.md 8(f) row 1: tests/wavefront_ref.py restates these functions in numpy (it has to: the image tests need
// the whole wavefront loop vectorised), and tests/test_shade_oracle.py holds every one of the restated functions against the reference's
// own code on random inputs.  What the host build needs is supplied here exactly as in ref_bsdf_driver.cpp: the reference's host float3
// math (pulled in through RenderCore_Bart/core_settings.h), empty shadows of the IntelliSense helpers (oracle/shim), a few CUDA
// intrinsics, and the __constant__ light tables of rendercore_optix7/kernels/.cuda.cu:24-28 as plain globals.
// Quirks of the reference that the driver makes deterministic (and the tests name):
//   * LightPickProb / RandomPointOnLight keep the potentials in float[MAXISLIGHTS = 8] and overrun it with more lights: ref_set_lights
//     refuses more than 8;
//   * RandomPointOnLight's point-light branch declares a LOCAL lightColor (:196) and leaves the output unset; pickProb is unset when no
//     light has potential: both are preset to NaN / 0 here.
#include "core_settings.h"
#include <cmath>
#include <cstring>
#include <vector>
#define __device__
#define __forceinline__ inline
#define LH2_DEVFUNC static inline
static inline float __uint_as_float( uint u ) { float f; memcpy( &f, &u, 4 ); return f; }
static inline uint __float_as_uint( float f ) { uint u; memcpy( &u, &f, 4 ); return u; }
static inline void __sincosf( float x, float* s, float* c ) { *s = sinf( x ), *c = cosf( x ); }
static uint skywidth, skyheight;   // globals SampleSkydome (tools_shared.h:183-190) refers to; not used here
static float3* skyPixels;
#define sqr sqr_shared             // lib/platform/system.h:106 already has a host sqr
#define adjoint false              // the parameter exists under nvcc only (lambert.h:73-75)
// the light tables of the cores (rendercore_optix7/kernels/.cuda.cu:24-28: __constant__ pointers + counts)
static CoreLightTri* areaLights;
static CorePointLight* pointLights;
static CoreSpotLight* spotLights;
static CoreDirectionalLight* directionalLights;
static int4 lightCounts;
#include "tools_shared.h"
#include "compatibility.h"         // lib/sharedBSDFs: REFERENCE_OF etc.; the cores include it in front of the BSDF they pick (disney.h:23)
#include "lambert.h"
#include "lights_shared.h"

namespace {
inline float3 f3( const float* p, int i ) { return make_float3( p[3 * i], p[3 * i + 1], p[3 * i + 2] ); }
inline void put( float* p, int i, const float3 v ) { p[3 * i] = v.x, p[3 * i + 1] = v.y, p[3 * i + 2] = v.z; }
std::vector<CoreLightTri> gArea;
std::vector<CorePointLight> gPoint;
std::vector<CoreSpotLight> gSpot;
std::vector<CoreDirectionalLight> gDir;
inline ShadingData shading( const float* color, const float* transmittance, const uint* params, int i )
{
	ShadingData sd;
	sd.color = f3( color, i ), sd.flags = 0, sd.transmittance = f3( transmittance, i ), sd.matID = 0;
	sd.tint = make_float4( 0 );
	sd.parameters = make_uint4( params[4 * i], params[4 * i + 1], params[4 * i + 2], params[4 * i + 3] );
	return sd;
}
}

extern "C" {

// params: the four parameter words per item (x: roughness in bits 24..31, z: transmission in bits 16..23, w: eta as float bits)
void ref_lambert_sample( int n, const float* color, const float* transmittance, const uint* params, const float* iN, const float* N, const float* wow,
	const float* distance, const float* r3, const float* r4, float* wiw, float* pdf, int* specular, float* value )
{
	for (int i = 0; i < n; i++)
	{
		const ShadingData sd = shading( color, transmittance, params, i );
		float3 R = make_float3( 0 );
		float p = 0;
		bool spec = false;
		const float3 v = SampleBSDF( sd, f3( iN, i ), f3( N, i ), make_float3( 0 ), f3( wow, i ), distance[i], r3[i], r4[i], R, p, spec );
		put( wiw, i, R ), put( value, i, v ), pdf[i] = p, specular[i] = spec ? 1 : 0;
	}
}

void ref_lambert_evaluate( int n, const float* color, const float* transmittance, const uint* params, const float* iN, const float* wow, const float* wiw,
	float* pdf, float* value )
{
	for (int i = 0; i < n; i++)
	{
		const ShadingData sd = shading( color, transmittance, params, i );
		float p = 0;
		const float3 v = EvaluateBSDF( sd, f3( iN, i ), make_float3( 0 ), f3( wow, i ), f3( wiw, i ), p );
		put( value, i, v ), pdf[i] = p;
	}
}

// the four arrays in the host layouts of CoreAPI_Base::SetLights (common_classes.h:275-345); returns non-zero for more lights than the
// reference's fixed potential[] array holds
int ref_set_lights( const void* area, int nArea, const void* point, int nPoint, const void* spot, int nSpot, const void* dir, int nDir )
{
	if (nArea + nPoint + nSpot + nDir > MAXISLIGHTS) return 1;
	gArea.assign( (const CoreLightTri*)area, (const CoreLightTri*)area + nArea );
	gPoint.assign( (const CorePointLight*)point, (const CorePointLight*)point + nPoint );
	gSpot.assign( (const CoreSpotLight*)spot, (const CoreSpotLight*)spot + nSpot );
	gDir.assign( (const CoreDirectionalLight*)dir, (const CoreDirectionalLight*)dir + nDir );
	areaLights = gArea.data(), pointLights = gPoint.data(), spotLights = gSpot.data(), directionalLights = gDir.data();
	lightCounts = make_int4( nArea, nPoint, nSpot, nDir );
	return 0;
}

void ref_random_barycentrics( int n, const float* r0, float* bary )
{
	for (int i = 0; i < n; i++) put( bary, i, RandomBarycentrics( r0[i] ) );
}

void ref_random_point_on_light( int n, const float* r0, const float* r1, const float* I, const float* N, float* P, float* pickProb, float* lightPdf, float* lightColor )
{
	for (int i = 0; i < n; i++)
	{
		float pick = 0, pdf = 0;
		float3 color = make_float3( NAN );
		const float3 p = RandomPointOnLight( r0[i], r1[i], f3( I, i ), f3( N, i ), pick, pdf, color );
		put( P, i, p ), put( lightColor, i, color ), pickProb[i] = pick, lightPdf[i] = pdf;
	}
}

// implicit-connection side of MIS: the probability that NEE at (O, N) would have picked area light idx, seen at point I (lights_shared.h:122-138)
void ref_light_pick_prob( int n, const int* idx, const float* O, const float* N, const float* I, float* prob )
{
	for (int i = 0; i < n; i++) prob[i] = LightPickProb( idx[i], f3( O, i ), f3( N, i ), f3( I, i ) );
}

void ref_calculate_light_pdf( int n, const float* D, const float* t, const float* area, const float* lightNormal, float* pdf )
{
	for (int i = 0; i < n; i++) pdf[i] = CalculateLightPDF( f3( D, i ), t[i], area[i], f3( lightNormal, i ) );
}

void ref_safe_origin( int n, const float* O, const float* R, const float* N, float eps, float* out )
{
	for (int i = 0; i < n; i++) put( out, i, SafeOrigin( f3( O, i ), f3( R, i ), f3( N, i ), eps ) );
}

} // extern "C"
