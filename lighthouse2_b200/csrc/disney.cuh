// disney.cuh -- the Disney principled BSDF of the reference's GPU cores (Setting bsdf = 1), restated for the B200 shade stage.
// Follows lib/sharedBSDFs/disney.h (SampleBSDF :227-312, EvaluateBSDF :314-348, the component lobes :121-225) and
// lib/sharedBSDFs/ggxmdf.h (anisotropic GGX with visible-normal sampling :62-119, GTR1 clearcoat lobe :186-241), which the
// reference adapted from appleseed.  Parameters arrive as the reference keeps them: sixteen 0.8 fixed-point values in three words
// (rendercore_optix7/rendercore.cpp:511-535, tools_shared.h:40-54) plus eta as a float.  Quirks kept on purpose: the sheen lobe's
// half vector is normalize( wo + wo ) (disney.h:202), EvaluateBSDF is called with the unflipped shading normal (pathtracer.h:186),
// the dielectric branch is the pure-specular one of the Lambert shader (disney.h:238-263).  Included by wavefront.cu after V3.
#pragma once

namespace disney {

struct Data {
	V3 color, transmittance, tint;
	float luminance, metallic, subsurface, specular, roughness, specTint, anisotropic, sheen, sheenTint, clearcoat, clearcoatGloss, transmission, eta;
};

__device__ __forceinline__ float Byte( const uint32_t word, const int shift ) { return (float)((word >> shift) & 255u) * (1.0f / 255.0f); } // CHAR2FLT
__device__ __forceinline__ void Unpack( Data& d, const uint32_t px, const uint32_t py, const uint32_t pz )
{
	d.metallic = Byte( px, 0 ), d.subsurface = Byte( px, 8 ), d.specular = Byte( px, 16 ), d.roughness = fmaxf( 0.001f, Byte( px, 24 ) );
	d.specTint = Byte( py, 0 ), d.anisotropic = Byte( py, 8 ), d.sheen = Byte( py, 16 ), d.sheenTint = Byte( py, 24 );
	d.clearcoat = Byte( pz, 0 ), d.clearcoatGloss = Byte( pz, 8 ), d.transmission = Byte( pz, 16 );
}
// material_shared.h:19-35, 68-69: tint = the base colour at unit luminance, luminance = CIE Y
__device__ __forceinline__ void Tint( Data& d, const V3 base )
{
	const float X = fmaxf( 0.0f, 0.412453f * base.x + 0.357580f * base.y + 0.180423f * base.z );
	const float Y = fmaxf( 0.0f, 0.212671f * base.x + 0.715160f * base.y + 0.072169f * base.z );
	const float Z = fmaxf( 0.0f, 0.019334f * base.x + 0.119193f * base.y + 0.950227f * base.z );
	d.luminance = Y;
	if (Y > 0)
	{
		const float s = 1.0f / Y, x = X * s, y = Y * s, z = Z * s;
		d.tint = v3( fmaxf( 0.0f, 3.240479f * x - 1.537150f * y - 0.498535f * z ), fmaxf( 0.0f, -0.969256f * x + 1.875992f * y + 0.041556f * z ),
			fmaxf( 0.0f, 0.055648f * x - 0.204043f * y + 1.057311f * z ) );
	}
	else d.tint = v3( 1.f, 1.f, 1.f );
}

__device__ __forceinline__ float sqr( const float x ) { return x * x; }
__device__ __forceinline__ float lerp( const float a, const float b, const float t ) { return a + t * (b - a); }
__device__ __forceinline__ float mix( const float a, const float b, const float x ) { return x <= 0 ? a : x >= 1 ? b : lerp( a, b, x ); } // tools_shared.h:98
__device__ __forceinline__ V3 cross( const V3 a, const V3 b ) { return v3( a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x ); }
__device__ __forceinline__ V3 reflect( const V3 i, const V3 n ) { return i - n * 2.0f * dot( n, i ); }
__device__ __forceinline__ V3 ToLocal( const V3 v, const V3 N, const V3 T, const V3 B ) { return v3( dot( v, T ), dot( v, B ), dot( v, N ) ); } // World2Tangent
__device__ __forceinline__ V3 ToWorld( const V3 v, const V3 N, const V3 T, const V3 B ) { return T * v.x + B * v.y + N * v.z; }           // Tangent2World
__device__ __forceinline__ float SchlickFresnel( const float u )
{
	const float m = __saturatef( 1.0f - u ), m2 = m * m, m4 = m2 * m2;
	return m4 * m;
}
__device__ __forceinline__ void AlphaFromRoughness( const float roughness, const float anisotropy, float& ax, float& ay ) // disney.h:43-49
{
	const float sq = roughness * roughness, aspect = sqrtf( 1.0f + anisotropy * (anisotropy < 0 ? 0.9f : -0.9f) );
	ax = fmaxf( 0.001f, sq / aspect ), ay = fmaxf( 0.001f, sq * aspect );
}
__device__ __forceinline__ float ClearcoatRoughness( const Data& d ) { return mix( 0.1f, 0.001f, d.clearcoatGloss ); }

// ---- ggxmdf.h -------------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ float StretchedRoughness( const V3 m, const float sinTheta, const float ax, const float ay )
{
	if (ax == ay || sinTheta == 0.0f) return 1.0f / sqr( ax );
	return sqr( m.x / (sinTheta * ax) ) + sqr( m.y / (sinTheta * ay) );
}
__device__ __forceinline__ float ProjectedRoughness( const V3 m, const float sinTheta, const float ax, const float ay )
{
	if (ax == ay || sinTheta == 0.0f) return ax;
	return sqrtf( sqr( (m.x * ax) / sinTheta ) + sqr( (m.y * ay) / sinTheta ) );
}
__device__ __forceinline__ float GGX_D( const V3 m, const float ax, const float ay )
{
	const float c = m.z;
	if (c == 0.0f) return sqr( ax ) * INVPI;
	const float c2 = sqr( c ), sinTheta = sqrtf( fmaxf( 0.0f, 1.0f - c2 ) ), c4 = sqr( c2 ), tan2 = (1.0f - c2) / c2;
	const float tmp = 1.0f + tan2 * StretchedRoughness( m, sinTheta, ax, ay );
	return 1.0f / (PI * ax * ay * c4 * sqr( tmp ));
}
__device__ __forceinline__ float GGX_Lambda( const V3 v, const float ax, const float ay )
{
	const float c = v.z;
	if (c == 0.0f) return 0.0f;
	const float c2 = sqr( c ), sinTheta = sqrtf( fmaxf( 0.0f, 1.0f - c2 ) );
	const float alpha = ProjectedRoughness( v, sinTheta, ax, ay ), tan2 = sqr( sinTheta ) / c2;
	return (-1.0f + sqrtf( 1.0f + sqr( alpha ) * tan2 )) * 0.5f;
}
__device__ __forceinline__ float GGX_G( const V3 wi, const V3 wo, const float ax, const float ay ) { return 1.0f / (1.0f + GGX_Lambda( wo, ax, ay ) + GGX_Lambda( wi, ax, ay )); }
__device__ __forceinline__ float GGX_Pdf( const V3 v, const V3 m, const float ax, const float ay ) // pdf of the visible normals
{
	if (v.z == 0.0f) return 0;
	return (1.0f / (1.0f + GGX_Lambda( v, ax, ay ))) * fabsf( dot( v, m ) ) * GGX_D( m, ax, ay ) / fabsf( v.z );
}
__device__ __forceinline__ V3 GGX_Sample( const V3 v, const float r0, const float r1, const float ax, const float ay )
{
	const float sgn = v.z < 0.0f ? -1.0f : 1.0f;
	const V3 stretched = normalize( v3( sgn * v.x * ax, sgn * v.y * ay, sgn * v.z ) );
	const V3 t1 = v.z < 0.9999f ? normalize( cross( stretched, v3( 0.f, 0.f, 1.f ) ) ) : v3( 1.f, 0.f, 0.f );
	const V3 t2 = cross( t1, stretched );
	const float a = 1.0f / (1.0f + stretched.z), r = sqrtf( r0 );
	const float phi = r1 < a ? r1 / a * PI : PI + (r1 - a) / (1.0f - a) * PI;
	const float p1 = r * cosf( phi ), p2 = r * sinf( phi ) * (r1 < a ? 1.0f : stretched.z);
	const V3 h = t1 * p1 + t2 * p2 + stretched * sqrtf( fmaxf( 0.0f, 1.0f - p1 * p1 - p2 * p2 ) );
	return normalize( v3( h.x * ax, h.y * ay, fmaxf( 0.0f, h.z ) ) );
}
__device__ __forceinline__ float GTR1_D( const V3 m, const float ax )
{
	const float alpha = fminf( fmaxf( ax, 0.001f ), 0.999f ), a2 = sqr( alpha ), c2 = sqr( m.z );
	return ((a2 - 1.0f) / (PI * logf( a2 ))) * (1 / (1 + (a2 - 1) * c2));
}
__device__ __forceinline__ float GTR1_Lambda( const V3 v, const float ax )
{
	const float c = v.z;
	if (c == 0) return 0;
	const float c2 = sqr( c ), sinTheta = sqrtf( fmaxf( 0.0f, 1.0f - c2 ) );
	if (sinTheta == 0.0f) return 0.0f;
	const float cot2 = c2 / sqr( sinTheta ), cot = sqrtf( cot2 ), alpha = fminf( fmaxf( ax, 0.001f ), 0.999f ), a2 = sqr( alpha );
	const float a = sqrtf( cot2 + a2 ), b = sqrtf( cot2 + 1.0f ), cc = logf( cot + b ), dd = logf( cot + a );
	return (a - b + cot * (cc - dd)) / (cot * logf( a2 ));
}
__device__ __forceinline__ float GTR1_G( const V3 wi, const V3 wo, const float ax ) { return 1.0f / (1.0f + GTR1_Lambda( wo, ax ) + GTR1_Lambda( wi, ax )); }
__device__ __forceinline__ float GTR1_Pdf( const V3 m, const float ax ) { return GTR1_D( m, ax ) * fabsf( m.z ); }
__device__ __forceinline__ V3 GTR1_Sample( const float r0, const float r1, const float ax )
{
	const float alpha = fminf( fmaxf( ax, 0.001f ), 0.999f ), a2 = sqr( alpha );
	const float c2 = (1.0f - powf( a2, 1.0f - r0 )) / (1.0f - a2), c = sqrtf( c2 ), s = sqrtf( fmaxf( 0.0f, 1.0f - c2 ) );
	const float phi = TWOPI * r1;
	return v3( cosf( phi ) * s, sinf( phi ) * s, c );
}

// ---- disney.h: lobes --------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ V3 SpecularFresnel( const Data& d, const V3 o, const V3 h ) // :52-59
{
	V3 value = v3( (1.0f - d.specTint) + d.specTint * d.tint.x, (1.0f - d.specTint) + d.specTint * d.tint.y, (1.0f - d.specTint) + d.specTint * d.tint.z );
	value = value * (d.specular * 0.08f);
	value = value * (1.0f - d.metallic) + d.color * d.metallic;
	const float t = SchlickFresnel( fabsf( dot( o, h ) ) );
	return v3( (1.0f - t) * value.x + t, (1.0f - t) * value.y + t, (1.0f - t) * value.z + t );
}
__device__ __forceinline__ V3 ClearcoatFresnel( const Data& d, const V3 o, const V3 h ) // :60-64
{
	const float v = mix( 0.04f, 1.0f, SchlickFresnel( fabsf( dot( o, h ) ) ) ) * 0.25f * d.clearcoat;
	return v3( v, v, v );
}
__device__ __forceinline__ bool ForceAboveSurface( V3& direction, const V3 normal ) // :65-73
{
	const float correction = 1.0e-4f - dot( direction, normal );
	if (correction <= 0) return false;
	direction = normalize( direction + normal * correction );
	return true;
}
// :96-121; GGX = the specular lobe (anisotropic GGX), !GGX = the clearcoat lobe (GTR1)
template <bool GGX> __device__ __forceinline__ void SampleMicrofacet( const Data& d, const float r0, const float r1, const float ax, const float ay,
	const V3 N, const V3 T, const V3 B, const V3 gN, const V3 wow, V3& wiw, float& pdf, V3& value )
{
	const V3 wo = ToLocal( wow, N, T, B );
	if (wo.z == 0) return;
	V3 m = GGX ? GGX_Sample( wo, r0, r1, ax, ay ) : GTR1_Sample( r0, r1, ax );
	V3 wi = reflect( wo * -1.0f, m );
	if (ForceAboveSurface( wi, ToLocal( gN, N, T, B ) )) m = normalize( wo + wi ); // keep the sample above the geometric surface
	if (wi.z == 0) return;
	const float cosOH = dot( wo, m );
	pdf = (GGX ? GGX_Pdf( wo, m, ax, ay ) : GTR1_Pdf( m, ax )) / fabsf( 4.0f * cosOH );
	if (pdf < 1.0e-6f) return;
	const float D = GGX ? GGX_D( m, ax, ay ) : GTR1_D( m, ax ), G = GGX ? GGX_G( wi, wo, ax, ay ) : GTR1_G( wi, wo, ax );
	value = (GGX ? SpecularFresnel( d, wo, m ) : ClearcoatFresnel( d, wo, m )) * (D * G / fabsf( 4.0f * wo.z * wi.z ));
	wiw = ToWorld( wi, N, T, B );
}
// :123-140; returns the pdf
template <bool GGX> __device__ __forceinline__ float EvaluateMicrofacet( const Data& d, const float ax, const float ay, const V3 N, const V3 T, const V3 B,
	const V3 wow, const V3 wiw, V3& bsdf )
{
	bsdf = v3( 0.f, 0.f, 0.f );
	const V3 wo = ToLocal( wow, N, T, B ), wi = ToLocal( wiw, N, T, B );
	if (wo.z == 0 || wi.z == 0) return 0;
	const V3 m = normalize( wi + wo );
	const float cosOH = dot( wo, m );
	if (cosOH == 0) return 0;
	const float D = GGX ? GGX_D( m, ax, ay ) : GTR1_D( m, ax ), G = GGX ? GGX_G( wi, wo, ax, ay ) : GTR1_G( wi, wo, ax );
	bsdf = (GGX ? SpecularFresnel( d, wo, m ) : ClearcoatFresnel( d, wo, m )) * (D * G / fabsf( 4.0f * wo.z * wi.z ));
	return (GGX ? GGX_Pdf( wo, m, ax, ay ) : GTR1_Pdf( m, ax )) / fabsf( 4.0f * cosOH );
}
// :142-171; returns the pdf
__device__ __forceinline__ float EvaluateDiffuse( const Data& d, const V3 n, const V3 wow, const V3 wiw, V3& value )
{
	const V3 h = normalize( wiw + wow );
	const float cosON = dot( n, wow ), cosIN = dot( n, wiw ), cosIH = dot( wiw, h );
	const float fl = SchlickFresnel( cosIN ), fv = SchlickFresnel( cosON );
	float fd = 0;
	if (d.subsurface != 1.0f)
	{
		const float fd90 = 0.5f + 2.0f * sqr( cosIH ) * d.roughness;
		fd = mix( 1.0f, fd90, fl ) * mix( 1.0f, fd90, fv );
	}
	if (d.subsurface > 0)
	{
		// Hanrahan-Krueger approximation of an isotropic BSSRDF
		const float fss90 = sqr( cosIH ) * d.roughness, fss = mix( 1.0f, fss90, fl ) * mix( 1.0f, fss90, fv );
		const float ss = 1.25f * (fss * (1.0f / (fabsf( cosON ) + fabsf( cosIN )) - 0.5f) + 0.5f);
		fd = mix( fd, ss, d.subsurface );
	}
	value = d.color * fd * INVPI * (1.0f - d.metallic);
	return fabsf( cosIN ) * INVPI;
}
// :187-198; the half vector is normalize( wow + wow ) in the reference
__device__ __forceinline__ float EvaluateSheen( const Data& d, const V3 wow, const V3 wiw, V3& value )
{
	const V3 h = normalize( wow + wow );
	const float fh = SchlickFresnel( dot( wiw, h ) ), s = fh * d.sheen * (1.0f - d.metallic);
	value = v3( (1.0f - d.sheenTint) + d.sheenTint * d.tint.x, (1.0f - d.sheenTint) + d.sheenTint * d.tint.y, (1.0f - d.sheenTint) + d.sheenTint * d.tint.z ) * s;
	return 1.0f / (2 * PI);
}
__device__ __forceinline__ V3 CosineDirection( const float r0, const float r1, const V3 N, const V3 T, const V3 B ) // tools_shared.h:250-256
{
	const float term2 = sqrtf( 1 - r1 );
	float s, c;
	__sincosf( TWOPI * r0, &s, &c );
	return normalize( ToWorld( v3( c * term2, s * term2, sqrtf( r1 ) ), N, T, B ) );
}
__device__ __forceinline__ void Weights( const Data& d, float w[4] ) // :268-274
{
	w[0] = lerp( d.luminance, 0, d.metallic ), w[1] = lerp( d.sheen, 0, d.metallic ), w[2] = lerp( d.specular, 1, d.metallic ), w[3] = d.clearcoat * 0.25f;
	const float s = 1.0f / (w[0] + w[1] + w[2] + w[3]);
	w[0] *= s, w[1] *= s, w[2] *= s, w[3] *= s;
}

// disney.h:227-312.  iN: shading normal, N: geometric normal (both unflipped), iT: tangent, wow = -D.  Returns the BSDF value; pdf = 0 ends the path.
__device__ __forceinline__ V3 SampleBSDF( const Data& d, V3 iN, const V3 N, const V3 iT, const V3 wow, const float distance, const float r0, const float r1,
	V3& wiw, float& pdf, bool& specular )
{
	const float flip = dot( wow, N ) < 0 ? -1.f : 1.f;
	iN = iN * flip;
	wiw = iN, pdf = 0;
	if (r0 < d.transmission)
	{
		// the pure specular dielectric of the Lambert shader (:238-263)
		specular = true, pdf = 1;
		const float eio = flip < 0 ? 1.0f / d.eta : d.eta;
		float F = 1.0f;
		{
			float e = eio, c = dot( iN, wow ); // Fr_L (:74-83)
			if (c < 0.0f) e = 1.0f / e, c = fabsf( c );
			const float sinT2 = sqr( e ) * (1.0f - c * c);
			if (sinT2 <= 1.0f)
			{
				const float ldn = sqrtf( 1.0f - sinT2 ), ra = (c - e * ldn) / (c + e * ldn), rb = (ldn - e * c) / (ldn + e * c);
				F = 0.5f * (sqr( ra ) + sqr( rb ));
			}
		}
		const V3 beer = v3( expf( -d.transmittance.x * distance * 2.0f ), expf( -d.transmittance.y * distance * 2.0f ), expf( -d.transmittance.z * distance * 2.0f ) );
		if (r1 < F)
		{
			wiw = reflect( wow * -1.0f, iN );
			if (dot( N * flip, wiw ) <= 0) pdf = 0;
			return d.color * beer * (1 / fabsf( dot( iN, wiw ) ));
		}
		// Refract_L (:84-93)
		const float cosI = fabsf( dot( iN, wow ) ), sin2I = fmaxf( 0.0f, 1.0f - cosI * cosI ), sin2T = eio * eio * sin2I;
		if (sin2T >= 1) { pdf = 0; return v3( 0.f, 0.f, 0.f ); } // total internal reflection: the reference goes on with a zero throughput, the path ends here
		wiw = wow * -eio + iN * (eio * cosI - sqrtf( 1.0f - sin2T ));
		return d.color * beer * (1 / fabsf( dot( iN, wiw ) ));
	}
	const float r3 = (r0 - d.transmission) / (1 - d.transmission);
	const V3 B = normalize( cross( iN, iT ) ), T = normalize( cross( iN, B ) );
	float w[4];
	Weights( d, w );
	const float cdfX = w[0], cdfY = w[0] + w[1], cdfZ = w[0] + w[1] + w[2];
	float probability, componentPdf = 0;
	V3 contrib, value = v3( 0.f, 0.f, 0.f );
	if (r3 < cdfX)
	{
		wiw = CosineDirection( r3 / cdfX, r1, iN, T, B );
		componentPdf = EvaluateDiffuse( d, iN, wow, wiw, value );
		probability = w[0] * componentPdf, w[0] = 0;
	}
	else if (r3 < cdfY)
	{
		wiw = CosineDirection( (r3 - cdfX) / (cdfY - cdfX), r1, iN, T, B );
		componentPdf = EvaluateSheen( d, wow, wiw, value );
		probability = w[1] * componentPdf, w[1] = 0;
	}
	else if (r3 < cdfZ)
	{
		float ax, ay;
		AlphaFromRoughness( d.roughness, d.anisotropic, ax, ay );
		SampleMicrofacet<true>( d, (r3 - cdfY) / (cdfZ - cdfY), r1, ax, ay, iN, T, B, N * flip, wow, wiw, componentPdf, value );
		probability = w[2] * componentPdf, w[2] = 0;
	}
	else
	{
		const float alpha = ClearcoatRoughness( d );
		SampleMicrofacet<false>( d, (r3 - cdfZ) / (1 - cdfZ), r1, alpha, alpha, iN, T, B, N * flip, wow, wiw, componentPdf, value );
		probability = w[3] * componentPdf, w[3] = 0;
	}
	if (w[0] > 0) probability += w[0] * EvaluateDiffuse( d, iN, wow, wiw, contrib ), value = value + contrib;
	if (w[1] > 0) probability += w[1] * EvaluateSheen( d, wow, wiw, contrib ), value = value + contrib;
	if (w[2] > 0)
	{
		float ax, ay;
		AlphaFromRoughness( d.roughness, d.anisotropic, ax, ay );
		probability += w[2] * EvaluateMicrofacet<true>( d, ax, ay, iN, T, B, wow, wiw, contrib ), value = value + contrib;
	}
	if (w[3] > 0)
	{
		const float alpha = ClearcoatRoughness( d );
		probability += w[3] * EvaluateMicrofacet<false>( d, alpha, alpha, iN, T, B, wow, wiw, contrib ), value = value + contrib;
	}
	pdf = probability > 1.0e-6f ? probability : 0;
	return value;
}

// disney.h:314-348: value and pdf of the direction wiw, for next event estimation
__device__ __forceinline__ V3 EvaluateBSDF( const Data& d, const V3 iN, const V3 iT, const V3 wow, const V3 wiw, float& pdf )
{
	pdf = 0;
	V3 value = v3( 0.f, 0.f, 0.f ), contrib;
	if (d.transmission > 0.999f || d.roughness <= 0.001f) return value; // no explicit connections at specular vertices
	const V3 B = normalize( cross( iN, iT ) ), T = normalize( cross( iN, B ) );
	float w[4];
	Weights( d, w );
	// (each lobe overwrites `value` in the reference's first two calls, :329-330: the sheen value replaces the diffuse one)
	if (w[0] > 0) pdf += w[0] * EvaluateDiffuse( d, iN, wow, wiw, value );
	if (w[1] > 0) pdf += w[1] * EvaluateSheen( d, wow, wiw, value );
	if (w[2] > 0)
	{
		float ax, ay;
		AlphaFromRoughness( d.roughness, d.anisotropic, ax, ay );
		const float p = EvaluateMicrofacet<true>( d, ax, ay, iN, T, B, wow, wiw, contrib );
		if (p > 0) pdf += w[2] * p, value = value + contrib;
	}
	if (w[3] > 0)
	{
		const float alpha = ClearcoatRoughness( d );
		const float p = EvaluateMicrofacet<false>( d, alpha, alpha, iN, T, B, wow, wiw, contrib );
		if (p > 0) pdf += w[3] * p, value = value + contrib;
	}
	return value;
}

// "Consistent Normal Interpolation" (Reshetov et al. 2010) as in tools_shared.h:296-310: D points away from the surface
__device__ __forceinline__ V3 ConsistentNormal( const V3 D, const V3 iN, const float alpha )
{
	const float t = PI - 2 * alpha, q = (t * t) / (PI * (PI + (2 * PI - 4) * alpha));
	const float b = dot( D, iN ), g = 1 + q * (b - 1), rho = sqrtf( q * (1 + g) / (1 + b) );
	return normalize( D + (iN * (g + rho * b) - D * rho) );
}

} // namespace disney
