/* oracle/bart_port.c -- TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C restatement of the reference's CPU ray-traversal path (lib/RenderCore_Bart), used by
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs as the
 * CHECKER.  Nothing under lighthouse2_b200/ links, loads or calls this file; the product path has
 * no CPU fallback.
 *
 * Parity status: PINNED.  tests/test_oracle.py checks this port ray-for-ray (t bit-identical, prim
 * identical) against oracle/_ref/liboracle_ref.so, which is the reference's own unmodified
 * bvh.cpp/common.h compiled in place by oracle/Makefile, and against the golden hit buffers under
 * tests/golden/ that were generated from that library (tests/golden/make_golden.py).
 *
 * Every function cites the reference lines it follows.  The arithmetic is kept in the reference's
 * exact operation order (compile with -ffp-contract=off, no -ffast-math): float add/mul are never
 * fused, dot() is x*x+y*y+z*z left to right (common_types.h:349), cross() as common_types.h:388,
 * normalize() = v * (1.0f / sqrtf(dot(v,v))) (common_types.h:87,363), min/max are std::min/std::max
 * ((b<a)?b:a and (a<b)?b:a -- this fixes the NaN behaviour of the slab test).
 *
 * The port additionally records the instance index of the winning hit (the reference API returns
 * only material/N/t); test_oracle.py proves (t, prim) equal to the unmodified reference on every
 * ray, so the extra output does not alter the algorithm.
 */
#include <float.h>
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#define EPSILON 0.0001f /* lib/RenderSystem/common_settings.h:45 */

typedef struct { float x, y, z; } f3;
typedef struct { f3 min_bound; int first; f3 max_bound; int count; } Node; /* bvh.h:10-15, 32 bytes */

static inline f3 mk3( float x, float y, float z ) { f3 r = { x, y, z }; return r; }
static inline f3 sub3( f3 a, f3 b ) { return mk3( a.x - b.x, a.y - b.y, a.z - b.z ); }
static inline f3 add3( f3 a, f3 b ) { return mk3( a.x + b.x, a.y + b.y, a.z + b.z ); }
static inline f3 scale3( f3 a, float s ) { return mk3( a.x * s, a.y * s, a.z * s ); }
static inline float dot3( f3 a, f3 b ) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline f3 cross3( f3 a, f3 b ) { return mk3( a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x ); }
static inline float fminf_ref( float a, float b ) { return a < b ? a : b; } /* common_types.h:83 */
static inline float fmaxf_ref( float a, float b ) { return a > b ? a : b; } /* common_types.h:84 */
static inline f3 min3( f3 a, f3 b ) { return mk3( fminf_ref( a.x, b.x ), fminf_ref( a.y, b.y ), fminf_ref( a.z, b.z ) ); }
static inline f3 max3( f3 a, f3 b ) { return mk3( fmaxf_ref( a.x, b.x ), fmaxf_ref( a.y, b.y ), fmaxf_ref( a.z, b.z ) ); }
static inline float stdmin( float a, float b ) { return (b < a) ? b : a; }
static inline float stdmax( float a, float b ) { return (a < b) ? b : a; }
static inline f3 normalize3( f3 v ) { float invLen = 1.0f / sqrtf( dot3( v, v ) ); return scale3( v, invLen ); }
static inline float get3( f3 v, int i ) { return i == 0 ? v.x : (i == 1 ? v.y : v.z); }

/* mat4::TransformVector / TransformPoint, common_types.h:616-635 (row-major cell[16]) */
static inline f3 xform_vector( const float* c, f3 v )
{
	return mk3( c[0] * v.x + c[1] * v.y + c[2] * v.z, c[4] * v.x + c[5] * v.y + c[6] * v.z, c[8] * v.x + c[9] * v.y + c[10] * v.z );
}
static inline f3 xform_point( const float* c, f3 v )
{
	f3 res = mk3( c[0] * v.x + c[1] * v.y + c[2] * v.z + c[3], c[4] * v.x + c[5] * v.y + c[6] * v.z + c[7],
		c[8] * v.x + c[9] * v.y + c[10] * v.z + c[11] );
	float w = c[12] * v.x + c[13] * v.y + c[14] * v.z + c[15];
	if (w == 1) return res;
	return scale3( res, 1.f / w );
}

/* mat4::Inverted, common_types.h:568-611 (cofactor expansion, then * 1/det; identity if det == 0) */
void port_mat4_inverted( const float* cell, float* out )
{
	const float inv[16] = {
		cell[5] * cell[10] * cell[15] - cell[5] * cell[11] * cell[14] - cell[9] * cell[6] * cell[15] +
		cell[9] * cell[7] * cell[14] + cell[13] * cell[6] * cell[11] - cell[13] * cell[7] * cell[10],
		-cell[1] * cell[10] * cell[15] + cell[1] * cell[11] * cell[14] + cell[9] * cell[2] * cell[15] -
		cell[9] * cell[3] * cell[14] - cell[13] * cell[2] * cell[11] + cell[13] * cell[3] * cell[10],
		cell[1] * cell[6] * cell[15] - cell[1] * cell[7] * cell[14] - cell[5] * cell[2] * cell[15] +
		cell[5] * cell[3] * cell[14] + cell[13] * cell[2] * cell[7] - cell[13] * cell[3] * cell[6],
		-cell[1] * cell[6] * cell[11] + cell[1] * cell[7] * cell[10] + cell[5] * cell[2] * cell[11] -
		cell[5] * cell[3] * cell[10] - cell[9] * cell[2] * cell[7] + cell[9] * cell[3] * cell[6],
		-cell[4] * cell[10] * cell[15] + cell[4] * cell[11] * cell[14] + cell[8] * cell[6] * cell[15] -
		cell[8] * cell[7] * cell[14] - cell[12] * cell[6] * cell[11] + cell[12] * cell[7] * cell[10],
		cell[0] * cell[10] * cell[15] - cell[0] * cell[11] * cell[14] - cell[8] * cell[2] * cell[15] +
		cell[8] * cell[3] * cell[14] + cell[12] * cell[2] * cell[11] - cell[12] * cell[3] * cell[10],
		-cell[0] * cell[6] * cell[15] + cell[0] * cell[7] * cell[14] + cell[4] * cell[2] * cell[15] -
		cell[4] * cell[3] * cell[14] - cell[12] * cell[2] * cell[7] + cell[12] * cell[3] * cell[6],
		cell[0] * cell[6] * cell[11] - cell[0] * cell[7] * cell[10] - cell[4] * cell[2] * cell[11] +
		cell[4] * cell[3] * cell[10] + cell[8] * cell[2] * cell[7] - cell[8] * cell[3] * cell[6],
		cell[4] * cell[9] * cell[15] - cell[4] * cell[11] * cell[13] - cell[8] * cell[5] * cell[15] +
		cell[8] * cell[7] * cell[13] + cell[12] * cell[5] * cell[11] - cell[12] * cell[7] * cell[9],
		-cell[0] * cell[9] * cell[15] + cell[0] * cell[11] * cell[13] + cell[8] * cell[1] * cell[15] -
		cell[8] * cell[3] * cell[13] - cell[12] * cell[1] * cell[11] + cell[12] * cell[3] * cell[9],
		cell[0] * cell[5] * cell[15] - cell[0] * cell[7] * cell[13] - cell[4] * cell[1] * cell[15] +
		cell[4] * cell[3] * cell[13] + cell[12] * cell[1] * cell[7] - cell[12] * cell[3] * cell[5],
		-cell[0] * cell[5] * cell[11] + cell[0] * cell[7] * cell[9] + cell[4] * cell[1] * cell[11] -
		cell[4] * cell[3] * cell[9] - cell[8] * cell[1] * cell[7] + cell[8] * cell[3] * cell[5],
		-cell[4] * cell[9] * cell[14] + cell[4] * cell[10] * cell[13] + cell[8] * cell[5] * cell[14] -
		cell[8] * cell[6] * cell[13] - cell[12] * cell[5] * cell[10] + cell[12] * cell[6] * cell[9],
		cell[0] * cell[9] * cell[14] - cell[0] * cell[10] * cell[13] - cell[8] * cell[1] * cell[14] +
		cell[8] * cell[2] * cell[13] + cell[12] * cell[1] * cell[10] - cell[12] * cell[2] * cell[9],
		-cell[0] * cell[5] * cell[14] + cell[0] * cell[6] * cell[13] + cell[4] * cell[1] * cell[14] -
		cell[4] * cell[2] * cell[13] - cell[12] * cell[1] * cell[6] + cell[12] * cell[2] * cell[5],
		cell[0] * cell[5] * cell[10] - cell[0] * cell[6] * cell[9] - cell[4] * cell[1] * cell[10] +
		cell[4] * cell[2] * cell[9] + cell[8] * cell[1] * cell[6] - cell[8] * cell[2] * cell[5] };
	const float det = cell[0] * inv[0] + cell[1] * inv[4] + cell[2] * inv[8] + cell[3] * inv[12];
	static const float ident[16] = { 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1 };
	memcpy( out, ident, 64 );
	if (det != 0)
	{
		const float invdet = 1.0f / det;
		for (int i = 0; i < 16; i++) out[i] = inv[i] * invdet;
	}
}

/* ---- Ray::IntersectTriangle, lib/RenderCore_Bart/common.h:19-50 ------------------------------- */
static inline int intersect_triangle( f3 O, f3 D, f3 vertex0, f3 vertex1, f3 vertex2, float* t )
{
	f3 edge1 = sub3( vertex1, vertex0 );
	f3 edge2 = sub3( vertex2, vertex0 );
	f3 h = cross3( D, edge2 );
	float a = dot3( edge1, h );
	if (a > -EPSILON && a < EPSILON) return 0;
	float f = (float)(1.0 / (double)a); /* common.h:31: f = 1.0 / a (double divide, rounded to float) */
	f3 s = sub3( O, vertex0 );
	float u = f * dot3( s, h );
	if (u < 0.0 || u > 1.0) return 0;
	f3 q = cross3( s, edge1 );
	float v = f * dot3( D, q );
	if (v < 0.0 || u + v > 1.0) return 0;
	float tt = f * dot3( edge2, q );
	if (tt > EPSILON && tt < 1 / EPSILON) { *t = tt; return 1; }
	return 0;
}

/* ---- BVH::IntersectAABB, lib/RenderCore_Bart/bvh.cpp:7-42 -------------------------------------- */
static inline int intersect_aabb( f3 O, f3 D, const Node* n, float* t )
{
	float dfx = 1.0f / D.x, dfy = 1.0f / D.y, dfz = 1.0f / D.z;
	float t1 = (n->min_bound.x - O.x) * dfx;
	float t2 = (n->max_bound.x - O.x) * dfx;
	float t3 = (n->min_bound.y - O.y) * dfy;
	float t4 = (n->max_bound.y - O.y) * dfy;
	float t5 = (n->min_bound.z - O.z) * dfz;
	float t6 = (n->max_bound.z - O.z) * dfz;
	float t_min = stdmax( stdmax( stdmin( t1, t2 ), stdmin( t3, t4 ) ), stdmin( t5, t6 ) );
	float t_max = stdmin( stdmin( stdmax( t1, t2 ), stdmax( t3, t4 ) ), stdmax( t5, t6 ) );
	if (t_max < 0) { *t = t_max; return 0; }
	if (t_min > t_max) { *t = t_max; return 0; }
	*t = t_min;
	return 1;
}

/* ---- scene containers: Mesh (common.h:53-79), Instance (common.h:81-87) ------------------------ */
typedef struct {
	int N, pool_pointer;
	Node* pool;
	int* tri_indices;
	int *left, *right, *indices; /* scratch replacing the per-node new[] of bvh.cpp:97-98,183 */
} Bvh;

typedef struct {
	int triCount;
	f3 *v0, *v1, *v2, *centers, *minb, *maxb;
	f3 aabb_min, aabb_max, aabb_center;
	Bvh bvh;
} Mesh;

typedef struct { int mesh; float transform[16], inv_transform[16]; } Instance;

typedef struct {
	Mesh** meshes; int meshCount, meshCap;
	Instance* instances; int instCount, instCap;
	Bvh top; f3* topCenters; /* transformed mesh-AABB centres, recomputed per Rebuild */
	double blasBuildSeconds, tlasBuildSeconds;
} Scene;

static double now_s( void ) { struct timespec ts; clock_gettime( CLOCK_MONOTONIC, &ts ); return ts.tv_sec + 1e-9 * ts.tv_nsec; }

/* The BLAS and TLAS builders of the reference are the same code over different primitives
 * (bvh.cpp:57-256 vs :373-592).  `prim bounds for SplitCost`, `centre`, are abstracted here:
 *   BLAS: SplitCost over triangle vertex bounds (bvh.cpp:76-94), centres = tri_centers (:111-114)
 *   TLAS: SplitCost over UNTRANSFORMED mesh root bounds (:398-414), centres = transform.TransformPoint(aabb_center) (:428-434) */
typedef struct { const f3 *minb, *maxb, *centers; } PrimView;

static float split_cost( const PrimView* pv, const int* indices, int first, int count )
{
	f3 min_b = mk3( FLT_MAX, FLT_MAX, FLT_MAX ), max_b = mk3( -FLT_MAX, -FLT_MAX, -FLT_MAX );
	for (int i = first; i < first + count; i++)
	{
		min_b = min3( min_b, pv->minb[indices[i]] );
		max_b = max3( max_b, pv->maxb[indices[i]] );
	}
	f3 dims = sub3( max_b, min_b );
	float area = 2 * dims.x * dims.y + 2 * dims.x * dims.z + 2 * dims.y * dims.z;
	return count * area; /* count==0 -> 0*inf = NaN, which loses every comparison below, as in the reference */
}

/* BVH2::Partition bvh.cpp:96-178 / TopLevelBVH::Partition :416-507 */
static int partition( Bvh* b, const PrimView* pv, Node* node, int* indices, int* counts )
{
	const int bins = 8;
	int *left = b->left, *right = b->right;
	int* tri_indices = b->tri_indices;
	float base_cost = split_cost( pv, tri_indices, node->first, node->count );
	float best_split_cost = FLT_MAX, best_split_pos = 0;
	int best_split_plane = 0;
	f3 c_min = mk3( FLT_MAX, FLT_MAX, FLT_MAX ), c_max = mk3( -FLT_MAX, -FLT_MAX, -FLT_MAX );
	for (int i = node->first; i < node->first + node->count; i++)
	{
		c_min = min3( c_min, pv->centers[tri_indices[i]] );
		c_max = max3( c_max, pv->centers[tri_indices[i]] );
	}
	for (int i = 0; i < 3; i++)
	{
		if (get3( c_min, i ) == get3( c_max, i )) continue;
		float interval = (get3( c_max, i ) - get3( c_min, i )) / bins;
		for (int j = 0; j < bins; j++)
		{
			float pos = get3( c_min, i ) + j * interval;
			counts[0] = 0, counts[1] = 0;
			for (int k = node->first; k < node->first + node->count; k++)
			{
				if (get3( pv->centers[tri_indices[k]], i ) <= pos) left[counts[0]++] = tri_indices[k];
				else right[counts[1]++] = tri_indices[k];
			}
			float sc = split_cost( pv, left, 0, counts[0] ) + split_cost( pv, right, 0, counts[1] );
			if (sc + EPSILON < best_split_cost) { best_split_cost = sc; best_split_plane = i; best_split_pos = pos; }
		}
	}
	if (base_cost - EPSILON < best_split_cost) return 0;
	counts[0] = 0, counts[1] = 0;
	for (int i = node->first; i < node->first + node->count; i++)
	{
		if (get3( pv->centers[tri_indices[i]], best_split_plane ) <= best_split_pos) left[counts[0]++] = tri_indices[i];
		else right[counts[1]++] = tri_indices[i];
	}
	for (int i = 0; i < counts[0]; i++) indices[i] = left[i];
	for (int i = 0; i < counts[1]; i++) indices[counts[0] + i] = right[i];
	return 1;
}

/* BVH2::Subdivide + SubdivideRecursively, bvh.cpp:180-224 (TLAS: :509-553) */
static void subdivide_recursively( Bvh* b, const PrimView* pv, int nodeIdx )
{
	Node* node = &b->pool[nodeIdx];
	if (node->count < 4) return;
	int counts[2] = { 0, 0 };
	/* `indices` scratch is consumed before recursing, so one shared buffer offset by node->first is safe */
	int* indices = b->indices + node->first;
	if (!partition( b, pv, node, indices, counts )) return;
	for (int i = 0; i < node->count; i++) b->tri_indices[node->first + i] = indices[i];
	int l = b->pool_pointer, r = b->pool_pointer + 1;
	b->pool[l].first = node->first, b->pool[l].count = counts[0];
	b->pool[r].first = node->first + counts[0], b->pool[r].count = counts[1];
	node->count = 0;
	node->first = b->pool_pointer;
	b->pool_pointer += 2;
	subdivide_recursively( b, pv, l );
	subdivide_recursively( b, pv, r );
}

static void bvh_alloc( Bvh* b, int N )
{
	free( b->pool ), free( b->tri_indices ), free( b->left ), free( b->right ), free( b->indices );
	b->N = N;
	b->tri_indices = (int*)malloc( sizeof( int ) * (N > 0 ? N : 1) );
	for (int i = 0; i < N; i++) b->tri_indices[i] = i;
	b->pool = (Node*)malloc( sizeof( Node ) * (N > 0 ? 2 * N : 2) );
	for (int i = 0; i < 2 * N; i++) b->pool[i].count = -1; /* BVHNode::count default, bvh.h:14 */
	b->left = (int*)malloc( sizeof( int ) * (N > 0 ? N : 1) );
	b->right = (int*)malloc( sizeof( int ) * (N > 0 ? N : 1) );
	b->indices = (int*)malloc( sizeof( int ) * (N > 0 ? N : 1) );
	b->pool_pointer = 2; /* slot 1 stays unused so sibling pairs share a cache line, bvh.cpp:65-67 */
	b->pool[0].first = 0;
	b->pool[0].count = N;
}

/* BVH2::Rebuild bvh.cpp:57-74 + UpdateBounds :226-256 */
static void blas_rebuild( Mesh* m )
{
	Bvh* b = &m->bvh;
	bvh_alloc( b, m->triCount );
	PrimView pv = { m->minb, m->maxb, m->centers };
	subdivide_recursively( b, &pv, 0 );
	for (int i = b->pool_pointer - 1; i >= 0; i--)
	{
		Node* n = &b->pool[i];
		if (n->count == -1) continue;
		if (n->count == 0)
		{
			const Node *l = &b->pool[n->first], *r = &b->pool[n->first + 1];
			f3 mn = min3( l->min_bound, r->min_bound ), mx = max3( l->max_bound, r->max_bound );
			n->min_bound = mn, n->max_bound = mx;
			continue;
		}
		n->min_bound = mk3( FLT_MAX, FLT_MAX, FLT_MAX ), n->max_bound = mk3( -FLT_MAX, -FLT_MAX, -FLT_MAX );
		for (int j = n->first; j < n->first + n->count; j++)
		{
			n->min_bound = min3( n->min_bound, m->minb[b->tri_indices[j]] );
			n->max_bound = max3( n->max_bound, m->maxb[b->tri_indices[j]] );
		}
	}
}

/* TopLevelBVH::Rebuild bvh.cpp:373-396 + UpdateBounds :555-592 */
static void tlas_rebuild( Scene* s )
{
	Bvh* b = &s->top;
	int N = s->instCount;
	bvh_alloc( b, N );
	free( s->topCenters );
	s->topCenters = (f3*)malloc( sizeof( f3 ) * (N > 0 ? N : 1) );
	f3* rootMin = (f3*)malloc( sizeof( f3 ) * (N > 0 ? N : 1) );
	f3* rootMax = (f3*)malloc( sizeof( f3 ) * (N > 0 ? N : 1) );
	for (int i = 0; i < N; i++)
	{
		const Mesh* m = s->meshes[s->instances[i].mesh];
		s->topCenters[i] = xform_point( s->instances[i].transform, m->aabb_center );
		rootMin[i] = m->bvh.pool[0].min_bound, rootMax[i] = m->bvh.pool[0].max_bound;
	}
	PrimView pv = { rootMin, rootMax, s->topCenters };
	subdivide_recursively( b, &pv, 0 );
	for (int i = b->pool_pointer - 1; i >= 0; i--)
	{
		Node* n = &b->pool[i];
		if (n->count == -1) continue;
		if (n->count == 0)
		{
			const Node *l = &b->pool[n->first], *r = &b->pool[n->first + 1];
			f3 mn = min3( l->min_bound, r->min_bound ), mx = max3( l->max_bound, r->max_bound );
			n->min_bound = mn, n->max_bound = mx;
			continue;
		}
		f3 min_b = mk3( FLT_MAX, FLT_MAX, FLT_MAX ), max_b = mk3( -FLT_MAX, -FLT_MAX, -FLT_MAX );
		for (int j = n->first; j < n->first + n->count; j++)
		{
			const Instance* in = &s->instances[b->tri_indices[j]];
			const Mesh* m = s->meshes[in->mesh];
			/* only the two extreme corners are transformed (bvh.cpp:579-587): exact for translation */
			f3 mesh_minb = xform_point( in->transform, m->aabb_min );
			f3 mesh_maxb = xform_point( in->transform, m->aabb_max );
			min_b = min3( min_b, mesh_minb );
			max_b = max3( max_b, mesh_maxb );
		}
		n->min_bound = min_b, n->max_bound = max_b;
	}
	free( rootMin ), free( rootMax );
}

/* ---- BVH2::Traverse, bvh.cpp:258-302 ------------------------------------------------------------ */
static int blas_traverse( const Mesh* m, f3 O, f3 D, int* prim, float* t, int pool_index, int* visits )
{
	if (visits) (*visits)++;
	const Bvh* b = &m->bvh;
	const Node* n = &b->pool[pool_index];
	float initial_t = *t, found_t = *t;
	if (n->count > 0)
	{
		for (int i = n->first; i < n->first + n->count; i++)
		{
			int ti = b->tri_indices[i];
			if (intersect_triangle( O, D, m->v0[ti], m->v1[ti], m->v2[ti], &found_t ) && found_t < *t)
			{
				*t = found_t;
				*prim = ti;
			}
		}
	}
	else
	{
		float t_left = FLT_MAX, t_right = FLT_MAX;
		if (intersect_aabb( O, D, &b->pool[n->first], &t_left ))
		{
			if (intersect_aabb( O, D, &b->pool[n->first + 1], &t_right ))
			{
				if (t_left < t_right)
				{
					blas_traverse( m, O, D, prim, t, n->first, visits );
					blas_traverse( m, O, D, prim, t, n->first + 1, visits );
				}
				else
				{
					blas_traverse( m, O, D, prim, t, n->first + 1, visits );
					blas_traverse( m, O, D, prim, t, n->first, visits );
				}
			}
			else blas_traverse( m, O, D, prim, t, n->first, visits );
		}
		else if (intersect_aabb( O, D, &b->pool[n->first + 1], &t_right ))
			blas_traverse( m, O, D, prim, t, n->first + 1, visits );
	}
	return *t < initial_t;
}

/* ---- TopLevelBVH::Traverse, bvh.cpp:594-649 ------------------------------------------------------ */
static int tlas_traverse( const Scene* s, f3 O, f3 D, int* prim, int* inst, float* t, int pool_index, int* visits )
{
	if (visits) (*visits)++;
	const Bvh* b = &s->top;
	const Node* n = &b->pool[pool_index];
	float initial_t = *t, found_t = *t;
	if (n->count > 0)
	{
		for (int i = n->first; i < n->first + n->count; i++)
		{
			float t_aabb = FLT_MAX;
			const Instance* in = &s->instances[b->tri_indices[i]];
			const Mesh* m = s->meshes[in->mesh];
			f3 O2 = xform_point( in->inv_transform, O );
			f3 D2 = normalize3( xform_vector( in->inv_transform, D ) );
			if (intersect_aabb( O2, D2, &m->bvh.pool[0], &t_aabb ) && t_aabb < *t)
			{
				blas_traverse( m, O2, D2, prim, t, 0, visits );
				if (*t < found_t)
				{
					found_t = *t;
					*inst = b->tri_indices[i]; /* extra output; the reference updates N here (bvh.cpp:616-619) */
				}
			}
		}
	}
	else
	{
		float t_left = FLT_MAX, t_right = FLT_MAX;
		if (intersect_aabb( O, D, &b->pool[n->first], &t_left ) && t_left < *t)
		{
			if (intersect_aabb( O, D, &b->pool[n->first + 1], &t_right ) && t_right < *t)
			{
				if (t_left < t_right)
				{
					tlas_traverse( s, O, D, prim, inst, t, n->first, visits );
					tlas_traverse( s, O, D, prim, inst, t, n->first + 1, visits );
				}
				else
				{
					tlas_traverse( s, O, D, prim, inst, t, n->first + 1, visits );
					tlas_traverse( s, O, D, prim, inst, t, n->first, visits );
				}
			}
			else tlas_traverse( s, O, D, prim, inst, t, n->first, visits );
		}
		else if (intersect_aabb( O, D, &b->pool[n->first + 1], &t_right ) && t_right < *t)
			tlas_traverse( s, O, D, prim, inst, t, n->first + 1, visits );
	}
	return *t < initial_t;
}

/* ======================================= C ABI =================================================== */

void* port_scene_create( void ) { return calloc( 1, sizeof( Scene ) ); }

static void bvh_free( Bvh* b ) { free( b->pool ), free( b->tri_indices ), free( b->left ), free( b->right ), free( b->indices ); }
static void mesh_free( Mesh* m )
{
	free( m->v0 ), free( m->v1 ), free( m->v2 ), free( m->centers ), free( m->minb ), free( m->maxb );
	bvh_free( &m->bvh );
	free( m );
}
void port_scene_destroy( void* h )
{
	Scene* s = (Scene*)h;
	for (int i = 0; i < s->meshCount; i++) mesh_free( s->meshes[i] );
	free( s->meshes ), free( s->instances ), free( s->topCenters );
	bvh_free( &s->top );
	free( s );
}

/* RenderCore::SetGeometry, lib/RenderCore_Bart/rendercore.cpp:47-80.  verts: float[T][3][3]. */
int port_set_geometry( void* h, int meshIdx, const float* verts, int triCount )
{
	Scene* s = (Scene*)h;
	Mesh* m;
	if (triCount <= 0) return 2; /* the reference has undefined behaviour for empty meshes (new BVHNode[0], bvh.cpp:64-69) */
	if (meshIdx >= s->meshCount)
	{
		if (s->meshCount == s->meshCap)
		{
			s->meshCap = s->meshCap ? 2 * s->meshCap : 8;
			s->meshes = (Mesh**)realloc( s->meshes, sizeof( Mesh* ) * s->meshCap );
		}
		m = (Mesh*)calloc( 1, sizeof( Mesh ) );
		m->triCount = triCount;
		size_t nb = sizeof( f3 ) * (triCount > 0 ? triCount : 1);
		m->v0 = (f3*)malloc( nb ), m->v1 = (f3*)malloc( nb ), m->v2 = (f3*)malloc( nb );
		m->centers = (f3*)malloc( nb ), m->minb = (f3*)malloc( nb ), m->maxb = (f3*)malloc( nb );
		s->meshes[s->meshCount++] = m;
	}
	m = s->meshes[meshIdx]; /* existing mesh: triangle count assumed unchanged (rendercore.cpp:56) */
	if (triCount != m->triCount) return 1;
	m->aabb_min = mk3( FLT_MAX, FLT_MAX, FLT_MAX ), m->aabb_max = mk3( -FLT_MAX, -FLT_MAX, -FLT_MAX );
	for (int i = 0; i < triCount; i++)
	{
		m->v0[i] = mk3( verts[i * 9 + 0], verts[i * 9 + 1], verts[i * 9 + 2] );
		m->v1[i] = mk3( verts[i * 9 + 3], verts[i * 9 + 4], verts[i * 9 + 5] );
		m->v2[i] = mk3( verts[i * 9 + 6], verts[i * 9 + 7], verts[i * 9 + 8] );
		f3 sum = add3( add3( m->v0[i], m->v1[i] ), m->v2[i] );
		m->centers[i] = mk3( sum.x / 3.0f, sum.y / 3.0f, sum.z / 3.0f );
		m->minb[i] = min3( min3( m->v0[i], m->v1[i] ), m->v2[i] );
		m->maxb[i] = max3( max3( m->v0[i], m->v1[i] ), m->v2[i] );
		m->aabb_min = min3( m->aabb_min, m->minb[i] );
		m->aabb_max = max3( m->aabb_max, m->maxb[i] );
	}
	m->aabb_center = add3( m->aabb_min, scale3( sub3( m->aabb_max, m->aabb_min ), 0.5f ) );
	double t0 = now_s();
	blas_rebuild( m );
	s->blasBuildSeconds += now_s() - t0;
	return 0;
}

/* RenderCore::SetInstance, rendercore.cpp:82-101 (meshIdx == -1 truncates the instance list) */
int port_set_instance( void* h, int instIdx, int meshIdx, const float* mat16 )
{
	Scene* s = (Scene*)h;
	if (meshIdx == -1) { if (s->instCount > instIdx) s->instCount = instIdx; return 0; }
	if (meshIdx < 0 || meshIdx >= s->meshCount) return 1;
	if (instIdx >= s->instCount)
	{
		if (s->instCount == s->instCap)
		{
			s->instCap = s->instCap ? 2 * s->instCap : 16;
			s->instances = (Instance*)realloc( s->instances, sizeof( Instance ) * s->instCap );
		}
		instIdx = s->instCount++;
	}
	Instance* in = &s->instances[instIdx];
	in->mesh = meshIdx;
	memcpy( in->transform, mat16, 64 );
	port_mat4_inverted( mat16, in->inv_transform );
	return 0;
}

void port_build_toplevel( void* h )
{
	Scene* s = (Scene*)h;
	double t0 = now_s();
	tlas_rebuild( s );
	s->tlasBuildSeconds = now_s() - t0;
}

double port_blas_build_seconds( void* h ) { return ((Scene*)h)->blasBuildSeconds; }
double port_tlas_build_seconds( void* h ) { return ((Scene*)h)->tlasBuildSeconds; }
void port_get_inv_transform( void* h, int instIdx, float* out16 ) { memcpy( out16, ((Scene*)h)->instances[instIdx].inv_transform, 64 ); }

typedef struct {
	const Scene* s; int lo, hi; const float *O, *D; float* tOut; int *primOut, *instOut, *visitOut; int brute;
} TraceJob;

/* Exhaustive variant: every triangle of every instance with the reference's triangle arithmetic and
 * instance entry (bvh.cpp:607-610), no boxes.  Used by the tests to CLASSIFY disagreements: a hit the
 * exhaustive search finds but the reference BVH misses is a box-grazing ray (the reference's slab test
 * is not conservative), an equal-t different-prim pair is a tie. */
static void brute_one( const Scene* s, f3 O, f3 D, float* t, int* prim, int* inst )
{
	for (int k = 0; k < s->instCount; k++)
	{
		const Instance* in = &s->instances[k];
		const Mesh* m = s->meshes[in->mesh];
		f3 O2 = xform_point( in->inv_transform, O );
		f3 D2 = normalize3( xform_vector( in->inv_transform, D ) );
		for (int i = 0; i < m->triCount; i++)
		{
			float ft = *t;
			if (intersect_triangle( O2, D2, m->v0[i], m->v1[i], m->v2[i], &ft ) && ft < *t) *t = ft, *prim = i, *inst = k;
		}
	}
}

static void* trace_worker( void* arg )
{
	TraceJob* j = (TraceJob*)arg;
	for (int i = j->lo; i < j->hi; i++)
	{
		f3 o = mk3( j->O[i * 3], j->O[i * 3 + 1], j->O[i * 3 + 2] );
		f3 d = mk3( j->D[i * 3], j->D[i * 3 + 1], j->D[i * 3 + 2] );
		float t = FLT_MAX;
		int prim = -1, inst = -1, visits = 0;
		if (j->brute) brute_one( j->s, o, d, &t, &prim, &inst );
		else if (j->s->instCount > 0) tlas_traverse( j->s, o, d, &prim, &inst, &t, 0, j->visitOut ? &visits : 0 );
		int hit = t < FLT_MAX;
		j->tOut[i] = hit ? t : 1e34f;
		j->primOut[i] = hit ? prim : -1;
		if (j->instOut) j->instOut[i] = hit ? inst : -1;
		if (j->visitOut) j->visitOut[i] = visits;
	}
	return 0;
}

static void run_trace( void* h, int n, const float* O, const float* D, float* tOut, int* primOut, int* instOut, int* visitOut,
	int nthreads, int brute )
{
	if (nthreads < 1) nthreads = 1;
	if (nthreads > 256) nthreads = 256;
	pthread_t th[256];
	TraceJob jobs[256];
	int chunk = (n + nthreads - 1) / nthreads, used = 0;
	for (int k = 0; k < nthreads; k++)
	{
		int lo = k * chunk, hi = lo + chunk > n ? n : lo + chunk;
		if (lo >= hi) break;
		TraceJob jb = { (const Scene*)h, lo, hi, O, D, tOut, primOut, instOut, visitOut, brute };
		jobs[used] = jb;
		if (nthreads == 1) trace_worker( &jobs[used] );
		else pthread_create( &th[used], 0, trace_worker, &jobs[used] );
		used++;
	}
	if (nthreads > 1) for (int k = 0; k < used; k++) pthread_join( th[k], 0 );
}

/* O, D: float[n][3]; D is used verbatim (the caller applies the Ray ctor's normalize, port_normalize). */
void port_trace( void* h, int n, const float* O, const float* D, float* tOut, int* primOut, int* instOut, int nthreads )
{
	run_trace( h, n, O, D, tOut, primOut, instOut, 0, nthreads, 0 );
}
void port_trace_visits( void* h, int n, const float* O, const float* D, float* tOut, int* primOut, int* instOut, int* visitOut, int nthreads )
{
	run_trace( h, n, O, D, tOut, primOut, instOut, visitOut, nthreads, 0 );
}
void port_trace_brute( void* h, int n, const float* O, const float* D, float* tOut, int* primOut, int* instOut, int nthreads )
{
	run_trace( h, n, O, D, tOut, primOut, instOut, 0, nthreads, 1 );
}

/* t of ONE given (instance, prim) pair with the reference arithmetic; 1e34 if the test rejects. */
void port_trace_single( void* h, int n, const float* O, const float* D, const int* inst, const int* prim, float* tOut )
{
	const Scene* s = (const Scene*)h;
	for (int i = 0; i < n; i++)
	{
		tOut[i] = 1e34f;
		if (inst[i] < 0 || inst[i] >= s->instCount) continue;
		const Instance* in = &s->instances[inst[i]];
		const Mesh* m = s->meshes[in->mesh];
		if (prim[i] < 0 || prim[i] >= m->triCount) continue;
		f3 O2 = xform_point( in->inv_transform, mk3( O[i * 3], O[i * 3 + 1], O[i * 3 + 2] ) );
		f3 D2 = normalize3( xform_vector( in->inv_transform, mk3( D[i * 3], D[i * 3 + 1], D[i * 3 + 2] ) ) );
		float t = FLT_MAX;
		if (intersect_triangle( O2, D2, m->v0[prim[i]], m->v1[prim[i]], m->v2[prim[i]], &t )) tOut[i] = t;
	}
}

int port_intersect_triangle( const float* O, const float* D, const float* v9, float* tOut )
{
	return intersect_triangle( mk3( O[0], O[1], O[2] ), mk3( D[0], D[1], D[2] ), mk3( v9[0], v9[1], v9[2] ), mk3( v9[3], v9[4], v9[5] ),
		mk3( v9[6], v9[7], v9[8] ), tOut );
}
int port_intersect_aabb( const float* O, const float* D, const float* bmin, const float* bmax, float* tOut )
{
	Node n;
	n.min_bound = mk3( bmin[0], bmin[1], bmin[2] ), n.max_bound = mk3( bmax[0], bmax[1], bmax[2] );
	return intersect_aabb( mk3( O[0], O[1], O[2] ), mk3( D[0], D[1], D[2] ), &n, tOut );
}
void port_normalize( const float* v, float* out )
{
	f3 r = normalize3( mk3( v[0], v[1], v[2] ) );
	out[0] = r.x, out[1] = r.y, out[2] = r.z;
}
/* vectorised Ray-ctor normalisation for ray generators: D[n][3] in place */
void port_normalize_n( int n, float* D )
{
	for (int i = 0; i < n; i++)
	{
		f3 r = normalize3( mk3( D[i * 3], D[i * 3 + 1], D[i * 3 + 2] ) );
		D[i * 3] = r.x, D[i * 3 + 1] = r.y, D[i * 3 + 2] = r.z;
	}
}
/* tree statistics for the report */
int port_blas_node_count( void* h, int meshIdx ) { return ((Scene*)h)->meshes[meshIdx]->bvh.pool_pointer; }
