// oracle/ref_bsdf_driver.cpp -- TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// C-ABI array driver around the UNMODIFIED Disney BSDF of the reference's GPU cores, compiled for the host in place (never
// copied) from /root/reference by oracle/Makefile into oracle/_ref/liboracle_bsdf_ref.so:
//   SampleBSDF / EvaluateBSDF and the lobes    lib/sharedBSDFs/disney.h:121-348
//   GGX / GTR1 microfacet distributions        lib/sharedBSDFs/ggxmdf.h:22-241
//   ShadingData + parameter macros, mix, Tangent2World / World2Tangent, DiffuseReflectionCosWeighted, ConsistentNormal
//                                              lib/CUDA/shared_kernel_code/tools_shared.h:24-54, 98, 211-256, 296-310
// The headers are written for nvcc; what the host build needs is supplied here: the reference's own host float3 math
// (lib/RenderSystem/common_types.h, pulled in through RenderCore_Bart/core_settings.h exactly as ref_driver.cpp does), empty
// shadows of the IntelliSense helpers noerrors.h / .cuda.h (oracle/shim), and the four CUDA intrinsics the code uses.
// disney.h leaves wiw / pdf unwritten on its early returns (:103, :110, :113, :254); the driver presets wiw = 0, pdf = 0 so that
// those cases are deterministic.
#include "core_settings.h"
#include <cmath>
#include <cstring>
#define __device__
#define __forceinline__ inline
#define LH2_DEVFUNC static inline
static inline float __uint_as_float( uint u ) { float f; memcpy( &f, &u, 4 ); return f; }
static inline uint __float_as_uint( float f ) { uint u; memcpy( &u, &f, 4 ); return u; }
static inline void __sincosf( float x, float* s, float* c ) { *s = sinf( x ), *c = cosf( x ); }
static uint skywidth, skyheight;   // globals SampleSkydome (tools_shared.h:183-190) refers to; not used here
static float3* skyPixels;
#define sqr sqr_shared             // lib/platform/system.h:106 already has a host sqr
#define adjoint false              // the parameter exists under nvcc only (disney.h:229-231)
#include "tools_shared.h"
#include "ggxmdf.h"
#include "disney.h"

namespace {
inline float3 f3( const float* p, int i ) { return make_float3( p[3 * i], p[3 * i + 1], p[3 * i + 2] ); }
inline void put( float* p, int i, const float3 v ) { p[3 * i] = v.x, p[3 * i + 1] = v.y, p[3 * i + 2] = v.z; }
inline ShadingData shading( const float* color, const float* transmittance, const float* tint4, const uint* params, int i )
{
	ShadingData sd;
	sd.color = f3( color, i ), sd.flags = 0, sd.transmittance = f3( transmittance, i ), sd.matID = 0;
	sd.tint = make_float4( tint4[4 * i], tint4[4 * i + 1], tint4[4 * i + 2], tint4[4 * i + 3] );
	sd.parameters = make_uint4( params[4 * i], params[4 * i + 1], params[4 * i + 2], params[4 * i + 3] );
	return sd;
}
}

extern "C" {

void ref_disney_sample( int n, const float* color, const float* transmittance, const float* tint4, const uint* params, const float* iN, const float* N,
	const float* iT, const float* wow, const float* distance, const float* r0, const float* r1, float* wiw, float* pdf, int* specular, float* value )
{
	for (int i = 0; i < n; i++)
	{
		const ShadingData sd = shading( color, transmittance, tint4, params, i );
		float3 R = make_float3( 0 );
		float p = 0;
		bool spec = false;
		const float3 v = SampleBSDF( sd, f3( iN, i ), f3( N, i ), f3( iT, i ), f3( wow, i ), distance[i], r0[i], r1[i], R, p, spec );
		put( wiw, i, R ), put( value, i, v ), pdf[i] = p, specular[i] = spec ? 1 : 0;
	}
}

void ref_disney_evaluate( int n, const float* color, const float* transmittance, const float* tint4, const uint* params, const float* iN, const float* iT,
	const float* wow, const float* wiw, float* pdf, float* value )
{
	for (int i = 0; i < n; i++)
	{
		const ShadingData sd = shading( color, transmittance, tint4, params, i );
		float p = 0;
		const float3 v = EvaluateBSDF( sd, f3( iN, i ), f3( iT, i ), f3( wow, i ), f3( wiw, i ), p );
		put( value, i, v ), pdf[i] = p;
	}
}

void ref_consistent_normal( int n, const float* D, const float* iN, const float* alpha, float* out )
{
	for (int i = 0; i < n; i++) put( out, i, ConsistentNormal( f3( D, i ), f3( iN, i ), alpha[i] ) );
}

} // extern "C"
