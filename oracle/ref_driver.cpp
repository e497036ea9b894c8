// oracle/ref_driver.cpp -- TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// Thin C-ABI driver around the UNMODIFIED reference CPU BVH core.  It is compiled together
// with /root/reference/lib/RenderCore_Bart/bvh.cpp (in place, never copied) by oracle/Makefile
// into oracle/_ref/liboracle_ref.so.  Everything that decides a hit is the reference's own code:
//   TopLevelBVH::Rebuild/Traverse   lib/RenderCore_Bart/bvh.cpp:373,594
//   BVH2::Rebuild/Traverse          lib/RenderCore_Bart/bvh.cpp:57,258
//   BVH::IntersectAABB              lib/RenderCore_Bart/bvh.cpp:7
//   Ray::IntersectTriangle          lib/RenderCore_Bart/common.h:19
// What is restated here (because rendercore.cpp does not compile on Linux: DWORD/GetTickCount/GL)
// is only the bookkeeping of RenderCore::SetGeometry / SetInstance
// (lib/RenderCore_Bart/rendercore.cpp:47-101): copy the CoreTris, per-triangle centroid/min/max,
// mesh AABB, BVH2 rebuild; Instance{mesh, M, M.Inverted()} appended to the top-level BVH.
//
// Primitive ids: Traverse returns only `material` (bvh.cpp:270-271), so the driver stores the
// triangle index in CoreTri.material (a free uint, common_classes.h:67).
#include "core_settings.h"
#include <thread>
#include <vector>
#include <chrono>

using namespace lh2core;

namespace {
struct RefScene {
	std::vector<Mesh*> meshes;
	std::vector<Instance*> instances;
	TopLevelBVH top;
	double blasBuildSeconds = 0, tlasBuildSeconds = 0;
};
double now() { return std::chrono::duration<double>( std::chrono::steady_clock::now().time_since_epoch() ).count(); }
}

extern "C" {

void* ref_scene_create() { return new RefScene(); }

// The reference never frees meshes either (rendercore.cpp keeps them for the process lifetime);
// we free what we own.
void ref_scene_destroy( void* h )
{
	RefScene* s = (RefScene*)h;
	for (auto i : s->instances) delete i;
	for (auto m : s->meshes) delete m;
	delete s;
}

// verts: float[triCount][3][3] (xyz of vertex0/1/2).  Mirrors RenderCore::SetGeometry, rendercore.cpp:47-80.
int ref_set_geometry( void* h, int meshIdx, const float* verts, int triCount )
{
	RefScene* s = (RefScene*)h;
	Mesh* mesh;
	if (triCount <= 0) return 2; // BVH2::Rebuild writes pool[0] of a zero-length array (bvh.cpp:64-69): undefined behaviour
	if (meshIdx >= (int)s->meshes.size())
	{
		s->meshes.push_back( mesh = new Mesh( triCount * 3, triCount ) );
		mesh->bvh = new BVH2( mesh );
	}
	mesh = s->meshes[meshIdx];
	mesh->aabb_min_bound = make_float3( FLT_MAX );
	mesh->aabb_max_bound = make_float3( -FLT_MAX );
	for (int i = 0; i < triCount; i++)
	{
		CoreTri tri;
		tri.vertex0 = make_float3( verts[i * 9 + 0], verts[i * 9 + 1], verts[i * 9 + 2] );
		tri.vertex1 = make_float3( verts[i * 9 + 3], verts[i * 9 + 4], verts[i * 9 + 5] );
		tri.vertex2 = make_float3( verts[i * 9 + 6], verts[i * 9 + 7], verts[i * 9 + 8] );
		tri.material = (uint)i; // prim id comes back through Traverse's `material`
		float3 N = cross( tri.vertex1 - tri.vertex0, tri.vertex2 - tri.vertex0 );
		float l = sqrtf( dot( N, N ) );
		if (l > 0) N = N * (1.0f / l);
		tri.Nx = N.x, tri.Ny = N.y, tri.Nz = N.z;
		mesh->triangles[i] = tri;
		mesh->tri_centers[i] = (tri.vertex0 + tri.vertex1 + tri.vertex2) / 3.0f;
		mesh->tri_min_bounds[i] = fminf( fminf( tri.vertex0, tri.vertex1 ), tri.vertex2 );
		mesh->tri_max_bounds[i] = fmaxf( fmaxf( tri.vertex0, tri.vertex1 ), tri.vertex2 );
		mesh->aabb_min_bound = fminf( mesh->aabb_min_bound, mesh->tri_min_bounds[i] );
		mesh->aabb_max_bound = fmaxf( mesh->aabb_max_bound, mesh->tri_max_bounds[i] );
	}
	mesh->aabb_center = mesh->aabb_min_bound + 0.5 * (mesh->aabb_max_bound - mesh->aabb_min_bound);
	double t0 = now();
	mesh->bvh->Rebuild();
	s->blasBuildSeconds += now() - t0;
	return 0;
}

// mat16: row-major 4x4 (mat4::cell).  Mirrors RenderCore::SetInstance, rendercore.cpp:82-101.
int ref_set_instance( void* h, int instIdx, int meshIdx, const float* mat16 )
{
	RefScene* s = (RefScene*)h;
	if (meshIdx < 0 || meshIdx >= (int)s->meshes.size()) return 1;
	mat4 M;
	memcpy( M.cell, mat16, 64 );
	if (instIdx >= (int)s->instances.size())
	{
		Instance* inst = new Instance( s->meshes[meshIdx], M );
		s->instances.push_back( inst );
		s->top.AddInstance( inst );
	}
	else
	{
		s->instances[instIdx]->mesh = s->meshes[meshIdx];
		s->instances[instIdx]->transform = M;
		s->instances[instIdx]->inv_transform = M.Inverted();
	}
	return 0;
}

// RenderCore::Render rebuilds the top level every frame (rendercore.cpp:199).
void ref_build_toplevel( void* h )
{
	RefScene* s = (RefScene*)h;
	double t0 = now();
	s->top.Rebuild();
	s->tlasBuildSeconds = now() - t0;
}

double ref_blas_build_seconds( void* h ) { return ((RefScene*)h)->blasBuildSeconds; }
double ref_tlas_build_seconds( void* h ) { return ((RefScene*)h)->tlasBuildSeconds; }

// Copy of the per-instance inverse transform exactly as the reference computed it (mat4::Inverted).
void ref_get_inv_transform( void* h, int instIdx, float* out16 )
{
	memcpy( out16, ((RefScene*)h)->instances[instIdx]->inv_transform.cell, 64 );
}

// O, D: float[n][3].  D is installed into Ray::D verbatim (the Ray ctor's normalize is bypassed by
// overwriting the public member), so the caller controls the exact float direction the reference
// traverses with.  Output: t (FLT_MAX -> 1e34f on miss) and prim (-1 on miss).
void ref_trace( void* h, int n, const float* O, const float* D, float* tOut, int* primOut, int nthreads )
{
	RefScene* s = (RefScene*)h;
	auto work = [=]( int lo, int hi ) {
		for (int i = lo; i < hi; i++)
		{
			float3 o = make_float3( O[i * 3], O[i * 3 + 1], O[i * 3 + 2] );
			float3 d = make_float3( D[i * 3], D[i * 3 + 1], D[i * 3 + 2] );
			Ray ray( o, make_float3( 0, 0, 1 ) );
			ray.D = d;
			int material = -1;
			float3 N = make_float3( 0 );
			float t = FLT_MAX;
			bool hit = s->top.Traverse( ray, material, N, t );
			tOut[i] = hit ? t : 1e34f;
			primOut[i] = hit ? material : -1;
		}
	};
	if (nthreads <= 1) { work( 0, n ); return; }
	std::vector<std::thread> pool;
	int chunk = (n + nthreads - 1) / nthreads;
	for (int k = 0; k < nthreads; k++)
	{
		int lo = k * chunk, hi = lo + chunk > n ? n : lo + chunk;
		if (lo < hi) pool.emplace_back( work, lo, hi );
	}
	for (auto& th : pool) th.join();
}

// Single primitive entry points for the known-answer tests.
int ref_intersect_triangle( const float* O, const float* D, const float* v9, float* tOut )
{
	Ray ray( make_float3( O[0], O[1], O[2] ), make_float3( 0, 0, 1 ) );
	ray.D = make_float3( D[0], D[1], D[2] );
	CoreTri tri;
	tri.vertex0 = make_float3( v9[0], v9[1], v9[2] );
	tri.vertex1 = make_float3( v9[3], v9[4], v9[5] );
	tri.vertex2 = make_float3( v9[6], v9[7], v9[8] );
	float t = *tOut;
	bool r = ray.IntersectTriangle( tri, t );
	*tOut = t;
	return r ? 1 : 0;
}

// IntersectAABB is protected in BVH; reach it through a trivial subclass (no reference code is changed).
struct AABBProbe : public TopLevelBVH {
	bool Probe( const Ray& r, const BVHNode& n, float& t ) { return IntersectAABB( r, n, t ); }
};
int ref_intersect_aabb( const float* O, const float* D, const float* bmin, const float* bmax, float* tOut )
{
	Ray ray( make_float3( O[0], O[1], O[2] ), make_float3( 0, 0, 1 ) );
	ray.D = make_float3( D[0], D[1], D[2] );
	BVHNode n;
	n.min_bound = make_float3( bmin[0], bmin[1], bmin[2] );
	n.max_bound = make_float3( bmax[0], bmax[1], bmax[2] );
	AABBProbe p;
	float t = *tOut;
	bool r = p.Probe( ray, n, t );
	*tOut = t;
	return r ? 1 : 0;
}

// The Ray ctor's normalisation (common.h:9-11 -> common_types.h:87,363), for ray generators.
void ref_normalize( const float* v, float* out )
{
	float3 r = normalize( make_float3( v[0], v[1], v[2] ) );
	out[0] = r.x, out[1] = r.y, out[2] = r.z;
}

} // extern "C"
