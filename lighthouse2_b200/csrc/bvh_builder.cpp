// bvh_builder.cpp -- host-side converter: primitive boxes -> binned-SAH BVH2 -> compressed 8-wide BVH.
//
// This is RenderCore_B200's own builder (it replaces BVH2::Rebuild, lib/RenderCore_Bart/bvh.cpp:57-256,
// and optixAccelBuild, lib/rendercore_optix7/core_mesh.cpp:94-130).  The tree it produces is NOT the
// reference's tree and does not need to be: the reference BLAS traversal never culls by distance
// (bvh.cpp:281-298), so any conservative hierarchy yields the same closest hit.
#include "wide_bvh.h"
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <emmintrin.h>
#include <memory>
#include <thread>

namespace b200 {

float g_sahPrimCost = 0.4f;
int g_sahLevels = 1;

namespace {

// primitive reference: two 16-byte halves, moved with aligned vector loads / stores; the centroid is recomputed where it is needed
struct alignas( 16 ) Ref { float lo[3]; uint32_t id; float hi[3]; uint32_t pad; };

struct Node2 {
	Aabb b;
	uint32_t left, right;   // children (inner) ...
	uint32_t first, count;  // ... or range in the ref array (leaf: count > 0)
	uint32_t span;          // primitives below this node (its range in the ref array starts at `first` for inner nodes too)
};

#ifndef B200_DP_LEAVES
#define B200_DP_LEAVES 1 // 1: the binary tree is split down to single primitives and the collapse decides which subtrees of <= maxLeaf primitives become leaves
#endif
#ifndef B200_SAH_BINS
#define B200_SAH_BINS 64 // 16 -> 64 bins: +2 % rays/s on the 1 M-triangle scene (primary +5 %), build 0.36 -> 0.40 s; 128..512 bins and an exact sweep of ranges below 16..64 primitives add nothing
#endif
constexpr int BINS = B200_SAH_BINS;
constexpr float COST_TRAVERSE = 1.0f;
#define COST_PRIM g_sahPrimCost // cost of a leaf triangle relative to a wide-node visit in the kernels (Setting "sahPrimCost")

} // namespace

// Frame of a node: per axis a power-of-two cell size s = 2^(e-127) and an origin p a little BELOW the box (s/8), chosen so that the box
// plus s/8 on either side spans at most 255 cells.  The headroom is what lets QuantiseChild keep its guard band on EVERY plane: with
// p = lo and the tightest s (the first version), the children that touch the node's own bounds -- there is one on every side -- got the
// planes 0 and 255 by clamping, i.e. no guard at all, and a ray through the very corner of such a box (a ray aimed at a mesh vertex)
// could lose the box to the rounding of the slab arithmetic (tests/test_fullsize_gpu.py::test_c2_exact_ties_do_not_depend_on_the_warp).
// refit.cu::frameOf restates this for the device; keep the two in step.
void QuantiseFrame( const Aabb& box, float p[3], uint8_t e[3] )
{
	for (int a = 0; a < 3; a++)
	{
		p[a] = box.lo[a];
		double ext = (double)box.hi[a] - (double)box.lo[a];
		int ex = -100; // 2^-100: harmless scale for flat axes (both planes of every child sit at p; the slab test's relative margins apply)
		if (ext > 0)
		{
			int fe;
			std::frexp( ext / 254.75, &fe ); // ext/254.75 = m * 2^fe, m in [0.5,1)  ->  254.75 * 2^fe >= ext
			ex = fe;
			if (std::ldexp( 254.75, ex - 1 ) >= ext) ex--; // m == 0.5 exactly: one less is still enough
			if (ex < -100) ex = -100;
			p[a] = (float)((double)box.lo[a] - std::ldexp( 1.0, ex - 3 )); // (cells below a few ulp of lo: rounds back to lo, the geometry is at float resolution there)
		}
		e[a] = (uint8_t)(ex + 127);
	}
}

void QuantiseChild( const float p[3], const uint8_t e[3], const Aabb& c, uint8_t qlo[3], uint8_t qhi[3] )
{
	const double guard = 1.0 / 64.0; // covers the 1/512-cell rounding of the one-FFMA plane decode (traverse.cu)
	for (int a = 0; a < 3; a++)
	{
		double inv = std::ldexp( 1.0, 127 - (int)e[a] );
		double lo = std::floor( ((double)c.lo[a] - (double)p[a]) * inv - guard );
		double hi = std::ceil( ((double)c.hi[a] - (double)p[a]) * inv + guard );
		qlo[a] = (uint8_t)(lo < 0 ? 0 : (lo > 255 ? 255 : lo));
		qhi[a] = (uint8_t)(hi < 0 ? 0 : (hi > 255 ? 255 : hi));
	}
}

} // namespace b200

// ---- BVH2 -> 8-wide collapse ------------------------------------------------------------------------
namespace b200 {

namespace {

// Fill slot metadata + quantised boxes of one wide node from its (already slot-ordered) children.
struct ChildDesc { Aabb b; bool inner; uint32_t primCount; uint32_t primOffset; };

void EncodeNode( WideNode& wn, const Aabb& box, const ChildDesc* slot[8] )
{
	memset( &wn.imask, 0, 1 );
	memset( wn.meta, 0, 8 );
	QuantiseFrame( box, wn.p, wn.e );
	wn.imask = 0;
	for (int s = 0; s < 8; s++)
	{
		uint8_t qlo[3] = { 0, 0, 0 }, qhi[3] = { 0, 0, 0 };
		if (slot[s])
		{
			QuantiseChild( wn.p, wn.e, slot[s]->b, qlo, qhi );
			if (slot[s]->inner) wn.imask |= (uint8_t)(1u << s), wn.meta[s] = (uint8_t)(0x20 | (24 + s));
			else wn.meta[s] = (uint8_t)((((1u << slot[s]->primCount) - 1u) << 5) | slot[s]->primOffset);
		}
		wn.qlox[s] = qlo[0], wn.qloy[s] = qlo[1], wn.qloz[s] = qlo[2];
		wn.qhix[s] = qhi[0], wn.qhiy[s] = qhi[1], wn.qhiz[s] = qhi[2];
	}
}

// Octant-aware slot assignment: slot s is visited first by rays whose direction signs are the
// complement of s's sign vector, so a child goes to the slot whose sign vector it is farthest
// AGAINST (minimal dot of (child centre - node centre) with the slot's direction).
void AssignSlots( const Aabb& box, const Aabb* childBox, int n, int slotOf[8] )
{
	float cost[8][8];
	float nc[3] = { 0.5f * (box.lo[0] + box.hi[0]), 0.5f * (box.lo[1] + box.hi[1]), 0.5f * (box.lo[2] + box.hi[2]) };
	for (int s = 0; s < 8; s++)
	{
		float ds[3] = { (s & 4) ? -1.f : 1.f, (s & 2) ? -1.f : 1.f, (s & 1) ? -1.f : 1.f };
		for (int i = 0; i < n; i++)
		{
			float d = 0;
			for (int a = 0; a < 3; a++) d += (0.5f * (childBox[i].lo[a] + childBox[i].hi[a]) - nc[a]) * ds[a];
			cost[s][i] = d;
		}
	}
	bool slotUsed[8] = { 0 }, childDone[8] = { 0 };
	for (int k = 0; k < n; k++)
	{
		float best = 3.0e38f;
		int bs = -1, bi = -1;
		for (int s = 0; s < 8; s++) if (!slotUsed[s])
			for (int i = 0; i < n; i++) if (!childDone[i] && cost[s][i] < best) best = cost[s][i], bs = s, bi = i;
		slotUsed[bs] = true, childDone[bi] = true, slotOf[bi] = bs;
	}
}

// SAH-optimal BVH2 -> 8-wide collapse by dynamic programming over the binary tree (after Ylitie, Karras,
// Laine: "Efficient Incoherent Ray Traversal on GPUs Through Compressed Wide BVHs", HPG 2017, sec. 4.1).
// cost[n][i-1] = cheapest way to represent the subtree of BVH2 node n as a forest of at most i roots
// (i = 1: a single wide node or leaf), with node cost = area * COST_TRAVERSE and leaf cost = area * prims *
// COST_PRIM.  A node's row needs the rows of its two children: the builder fills it when both subtrees are done.
struct CollapsePlan {
	// (arrays of the size of the node pool, left uninitialised: every row that is read has been written by node())
	std::unique_ptr<float[]> cost;      // [n*7 + (i-1)]
	std::unique_ptr<uint8_t[]> choice;  // [n*7 + (i-1)], i >= 2: 0 = "use i-1 roots", k>0 = left gets k roots, right i-k
	std::unique_ptr<uint8_t[]> split8;  // [n]: left's share of the 8 child slots when n becomes a wide node
	std::unique_ptr<uint8_t[]> asLeaf;  // [n]: an inner BVH2 node of <= maxLeaf primitives that is cheaper as ONE leaf (B200_DP_LEAVES)

	void allocate( size_t rows )
	{
		cost.reset( new float[rows * 7] ), choice.reset( new uint8_t[rows * 7] ), split8.reset( new uint8_t[rows] ), asLeaf.reset( new uint8_t[rows] );
	}
	void node( const Node2* N2, size_t n, int maxLeaf )
	{
		const Node2& nd = N2[n];
		float A = nd.b.halfArea();
		float* c = &cost[n * 7];
		split8[n] = 0, asLeaf[n] = 0;
		if (nd.count > 0) { for (int i = 0; i < 7; i++) c[i] = A * nd.count * COST_PRIM; return; }
		const float *cl = &cost[(size_t)nd.left * 7], *cr = &cost[(size_t)nd.right * 7];
		auto distribute = [&]( int j, int& bestK ) {
			float best = 3.0e38f;
			for (int k = 1; k < j; k++)
			{
				if (k > 7 || j - k > 7) continue;
				float v = cl[k - 1] + cr[j - k - 1];
				if (v < best) best = v, bestK = k;
			}
			return best;
		};
		int k8 = 4;
		c[0] = A * COST_TRAVERSE + distribute( 8, k8 );
		split8[n] = (uint8_t)k8;
		if (B200_DP_LEAVES && (int)nd.span <= maxLeaf && A * nd.span * COST_PRIM <= c[0]) c[0] = A * nd.span * COST_PRIM, asLeaf[n] = 1;
		choice[n * 7] = 0;
		for (int i = 2; i <= 7; i++)
		{
			int k = 1;
			float d = distribute( i, k );
			if (d < c[i - 2]) c[i - 1] = d, choice[n * 7 + i - 1] = (uint8_t)k;
			else c[i - 1] = c[i - 2], choice[n * 7 + i - 1] = 0;
		}
	}
	// the roots of the forest chosen for BVH2 node m when it may use up to i child slots
	void expand( const Node2* N2, uint32_t m, int i, uint32_t* out, int& cn ) const
	{
		const Node2& nd = N2[m];
		while (i > 1 && nd.count == 0 && choice[(size_t)m * 7 + i - 1] == 0) i--;
		if (nd.count > 0 || i == 1) { out[cn++] = m; return; }
		int k = choice[(size_t)m * 7 + i - 1];
		expand( N2, nd.left, k, out, cn );
		expand( N2, nd.right, i - k, out, cn );
	}
};

} // namespace

// ---- binned-SAH binary tree -------------------------------------------------------------------------------------------------------
// The tree is a function of the primitive boxes alone -- not of the thread count, the chunk sizes or the scheduling:
//   * every split is a STABLE partition (the two halves keep the order they had), written from one buffer to the other, so the order of
//     the primitives in a leaf is their input order;
//   * boxes, bin boxes and bin counts are min / max / integer sums, which do not depend on how the range was cut into chunks.
// The top of the tree (ranges of >= `big` references) is built level by level with every pass cut into chunks that all threads work
// on; below that the subtrees are independent tasks, largest first.  Per level a reference is read twice: once to bin it on the three
// axes at a time, once to move it, and the second pass also collects the boxes of the two halves.
namespace {

struct Box4 {
	__m128 lo, hi; // lanes 0..2; lane 3 is whatever the references carry there (index bits) and is never read
	void reset() { lo = _mm_set1_ps( 3.0e38f ), hi = _mm_set1_ps( -3.0e38f ); }
	void grow( const Box4& b ) { lo = _mm_min_ps( lo, b.lo ), hi = _mm_max_ps( hi, b.hi ); }
	void grow( __m128 l, __m128 h ) { lo = _mm_min_ps( lo, l ), hi = _mm_max_ps( hi, h ); }
	void growPoint( __m128 p ) { lo = _mm_min_ps( lo, p ), hi = _mm_max_ps( hi, p ); }
	Aabb aabb() const
	{
		alignas( 16 ) float l[4], h[4];
		_mm_store_ps( l, lo ), _mm_store_ps( h, hi );
		Aabb b;
		for (int a = 0; a < 3; a++) b.lo[a] = l[a], b.hi[a] = h[a];
		return b;
	}
	float halfArea() const // Aabb::halfArea: (x * y + y * z) + z * x, the same three products and two sums
	{
		const __m128 d = _mm_sub_ps( hi, lo ), m = _mm_mul_ps( d, _mm_shuffle_ps( d, d, _MM_SHUFFLE( 3, 0, 2, 1 ) ) ); // (xy, yz, zx, .)
		return _mm_cvtss_f32( _mm_add_ss( _mm_add_ss( m, _mm_shuffle_ps( m, m, 1 ) ), _mm_shuffle_ps( m, m, 2 ) ) );
	}
};

inline __m128 lanes012() { return _mm_castsi128_ps( _mm_set_epi32( 0, -1, -1, -1 ) ); }
// centroid as the first builder stored it: 0.5f * (lo + hi) per axis; lane 3 cleared first (index bits are denormals as floats)
inline __m128 centroidOf( __m128 lo, __m128 hi ) { return _mm_mul_ps( _mm_set1_ps( 0.5f ), _mm_add_ps( _mm_and_ps( lo, lanes012() ), _mm_and_ps( hi, lanes012() ) ) ); }

static_assert( BINS <= 64, "Bins::used is one 64-bit mask per axis" );
struct Bins {
	__m128 lo[3][BINS], hi[3][BINS];
	uint32_t count[3][BINS];
	uint64_t used[3]; // bins that hold something.  Most nodes are small: sweeping and clearing only these is what keeps them cheap
	Bins() { for (int a = 0; a < 3; a++) { used[a] = ~0ull; } clear(); }
	void clear() // back to empty: only what was touched
	{
		for (int a = 0; a < 3; a++)
		{
			for (uint64_t m = used[a]; m; m &= m - 1)
			{
				const int k = __builtin_ctzll( m );
				if (k < BINS) lo[a][k] = _mm_set1_ps( 3.0e38f ), hi[a][k] = _mm_set1_ps( -3.0e38f ), count[a][k] = 0;
			}
			used[a] = 0;
		}
	}
	void add( const Bins& o )
	{
		for (int a = 0; a < 3; a++)
		{
			for (uint64_t m = o.used[a]; m; m &= m - 1)
			{
				const int k = __builtin_ctzll( m );
				lo[a][k] = _mm_min_ps( lo[a][k], o.lo[a][k] ), hi[a][k] = _mm_max_ps( hi[a][k], o.hi[a][k] ), count[a][k] += o.count[a][k];
			}
			used[a] |= o.used[a];
		}
	}
};

// bin index of a centroid: (int)((c - lo) * scale) clamped to [0, nb-1], for the three axes at a time.  Clamping before the conversion
// gives what clamping the integer gave; an axis without extent has scale = inf and lands in bin 0 (it is never chosen).
struct BinMap {
	__m128 lo, scale, top;
	int nb;
	BinMap( const Box4& cbox, int bins ) : nb( bins )
	{
		alignas( 16 ) float l[4], h[4], s[4] = { 0, 0, 0, 0 };
		_mm_store_ps( l, cbox.lo ), _mm_store_ps( h, cbox.hi );
		for (int a = 0; a < 3; a++) s[a] = (float)nb / (h[a] - l[a]);
		l[3] = 0;
		lo = _mm_load_ps( l ), scale = _mm_load_ps( s ), top = _mm_set1_ps( (float)(nb - 1) );
	}
	void index( __m128 c, int k[4] ) const
	{
		__m128 x = _mm_mul_ps( _mm_sub_ps( c, lo ), scale );
		x = _mm_min_ps( _mm_max_ps( x, _mm_setzero_ps() ), top ); // max( NaN, 0 ) = 0
		_mm_store_si128( (__m128i*)k, _mm_cvttps_epi32( x ) );
	}
};

inline int binsFor( uint32_t count ) { return count >= 256 ? BINS : (BINS < 16 ? BINS : 16); } // the fine bins pay at the top of the tree, where the ranges are long

void binRange( const Ref* r, uint32_t count, const BinMap& map, Bins& B )
{
	for (uint32_t i = 0; i < count; i++)
	{
		const __m128 lo = _mm_load_ps( r[i].lo ), hi = _mm_load_ps( r[i].hi );
		alignas( 16 ) int k[4];
		map.index( centroidOf( lo, hi ), k );
		for (int a = 0; a < 3; a++)
			B.lo[a][k[a]] = _mm_min_ps( B.lo[a][k[a]], lo ), B.hi[a][k[a]] = _mm_max_ps( B.hi[a][k[a]], hi ), B.count[a][k[a]]++, B.used[a] |= 1ull << k[a];
	}
}

struct Split { int axis, bin; uint32_t leftCount; float cost; };

// Sweep of the bins: the cheapest of the (nb - 1) planes on each axis that has extent, the first one among equals.  Only a plane right
// behind a bin that holds something can be the first with its two sets, so the sweep visits those bins alone.
Split chooseSplit( const Bins& B, const Box4& cbox )
{
	Split best = { -1, 0, 0, 3.0e38f };
	const Aabb cb = cbox.aabb();
	for (int a = 0; a < 3; a++)
	{
		if (!(cb.hi[a] - cb.lo[a] > 0)) continue;
		int bin[BINS], bins = 0;
		for (uint64_t m = B.used[a]; m; m &= m - 1) bin[bins++] = __builtin_ctzll( m );
		float rightArea[BINS]; // [j]: the bins bin[j] .. bin[bins - 1] together
		Box4 acc;
		acc.reset();
		for (int j = bins - 1; j > 0; j--) acc.grow( B.lo[a][bin[j]], B.hi[a][bin[j]] ), rightArea[j] = acc.halfArea();
		acc.reset();
		uint32_t n = 0, total = 0;
		for (int j = 0; j < bins; j++) total += B.count[a][bin[j]];
		for (int j = 0; j + 1 < bins; j++)
		{
			acc.grow( B.lo[a][bin[j]], B.hi[a][bin[j]] );
			n += B.count[a][bin[j]];
			float cost = acc.halfArea() * n + rightArea[j + 1] * (total - n);
			if (cost < best.cost) best.cost = cost, best.axis = a, best.bin = bin[j], best.leftCount = n;
		}
	}
	return best;
}

struct HalfBoxes { Box4 box[2], cbox[2]; void reset() { box[0].reset(), box[1].reset(), cbox[0].reset(), cbox[1].reset(); } };

// stable split of a range: bins <= split.bin go to dstLeft, the others to dstRight, both in input order
void splitRange( const Ref* src, uint32_t count, const BinMap& map, const Split& split, Ref* dstLeft, Ref* dstRight, HalfBoxes& hb )
{
	Ref* dst[2] = { dstLeft, dstRight };
	for (uint32_t i = 0; i < count; i++)
	{
		const __m128 lo = _mm_load_ps( src[i].lo ), hi = _mm_load_ps( src[i].hi ), c = centroidOf( lo, hi );
		alignas( 16 ) int k[4];
		map.index( c, k );
		const int side = k[split.axis] > split.bin;
		hb.box[side].grow( lo, hi ), hb.cbox[side].growPoint( c );
		Ref* d = dst[side]++;
		_mm_store_ps( d->lo, lo ), _mm_store_ps( d->hi, hi );
	}
}

void boundsOf( const Ref* r, uint32_t count, Box4& box, Box4& cbox )
{
	for (uint32_t i = 0; i < count; i++)
	{
		const __m128 lo = _mm_load_ps( r[i].lo ), hi = _mm_load_ps( r[i].hi );
		box.grow( lo, hi ), cbox.growPoint( centroidOf( lo, hi ) );
	}
}

struct Builder2 {
	std::unique_ptr<Ref[]> buf[2];    // a split reads one and writes the other; leaves leave their references in buf[0]
	std::unique_ptr<Node2[]> nodes;   // 2n + 2, handed out in pairs
	std::atomic<uint32_t> nodeCount{ 0 };
	CollapsePlan plan;                // rows filled bottom-up as the subtrees complete
	int maxLeaf = 3, threads = 1;

	struct Work { uint32_t node, first, count; int src; Box4 box, cbox; };
	static Work work( uint32_t node, uint32_t first, uint32_t count, int src, const Box4& box, const Box4& cbox )
	{
		Work w;
		w.node = node, w.first = first, w.count = count, w.src = src, w.box = box, w.cbox = cbox;
		return w;
	}

	// what one thread carries through a subtree: bins (empty between uses) and a block of nodes of its own, so that the shared counter is
	// touched once per subtree, not once per node
	struct Local { Bins bins; uint32_t nextNode = 0; };
	uint32_t alloc2( Local* local )
	{
		if (!local) return nodeCount.fetch_add( 2 );
		local->nextNode += 2;
		return local->nextNode - 2;
	}

	void initNode( const Work& w )
	{
		Node2& node = nodes[w.node];
		node.b = w.box.aabb();
		node.first = w.first, node.count = w.count, node.span = w.count, node.left = node.right = 0;
	}
	// a leaf's references belong in buf[0]
	void settle( const Work& w )
	{
		if (w.src == 1) memcpy( buf[0].get() + w.first, buf[1].get() + w.first, (size_t)w.count * sizeof( Ref ) );
	}
	// no plane separates the centroids: a leaf when small enough, else halves by position
	bool degenerate( const Work& w, Work child[2], Local* local )
	{
		if ((int)w.count <= maxLeaf) { settle( w ); return false; }
		const uint32_t lc = w.count / 2, l = alloc2( local );
		Box4 box[2], cbox[2];
		for (int s = 0; s < 2; s++)
			box[s].reset(), cbox[s].reset(), boundsOf( buf[w.src].get() + w.first + (s ? lc : 0), s ? w.count - lc : lc, box[s], cbox[s] );
		child[0] = work( l, w.first, lc, w.src, box[0], cbox[0] ), child[1] = work( l + 1, w.first + lc, w.count - lc, w.src, box[1], cbox[1] );
		nodes[w.node].left = l, nodes[w.node].right = l + 1, nodes[w.node].count = 0;
		return true;
	}
	void children( const Work& w, const Split& split, const HalfBoxes& hb, Work child[2], Local* local )
	{
		const uint32_t l = alloc2( local );
		child[0] = work( l, w.first, split.leftCount, w.src ^ 1, hb.box[0], hb.cbox[0] );
		child[1] = work( l + 1, w.first + split.leftCount, w.count - split.leftCount, w.src ^ 1, hb.box[1], hb.cbox[1] );
		nodes[w.node].left = l, nodes[w.node].right = l + 1, nodes[w.node].count = 0;
	}

	// one thread, one subtree, depth first; the plan rows follow on the way up
	void buildSubtree( const Work& w, Local& local )
	{
		initNode( w );
		if (w.count == 1) { settle( w ); plan.node( nodes.get(), w.node, maxLeaf ); return; }
		Work child[2];
		const BinMap map( w.cbox, binsFor( w.count ) );
		const Ref* src = buf[w.src].get() + w.first;
		binRange( src, w.count, map, local.bins );
		const Split split = chooseSplit( local.bins, w.cbox );
		local.bins.clear();
		if (split.axis >= 0)
		{
			HalfBoxes hb;
			hb.reset();
			Ref* dst = buf[w.src ^ 1].get() + w.first;
			splitRange( src, w.count, map, split, dst, dst + split.leftCount, hb );
			children( w, split, hb, child, &local );
		}
		else if (!degenerate( w, child, &local )) { plan.node( nodes.get(), w.node, maxLeaf ); return; }
		buildSubtree( child[0], local );
		buildSubtree( child[1], local );
		plan.node( nodes.get(), w.node, maxLeaf );
	}
	void buildTask( const Work& w )
	{
		std::unique_ptr<Local> local( new Local );
		local->nextNode = nodeCount.fetch_add( 2 * w.count - 2 ); // a tree over c references has 2c - 2 nodes below its root: 2n - 1 in all
		buildSubtree( w, *local );
	}

	void build( const Aabb* primBoxes, uint32_t n )
	{
		buf[0].reset( new Ref[n] ), buf[1].reset( new Ref[n] );
		nodes.reset( new Node2[2 * (size_t)n + 2] );
		plan.allocate( 2 * (size_t)n + 2 );
		nodeCount = 1;
		constexpr uint32_t CHUNK = 16384;
		// ---- references + root bounds
		const uint32_t rootChunks = (n + CHUNK - 1) / CHUNK;
		std::vector<Box4> part( 2 * (size_t)rootChunks );
		ParallelFor( threads, rootChunks, [&]( size_t c ) {
			Box4 box, cbox;
			box.reset(), cbox.reset();
			const uint32_t end = (uint32_t)std::min<uint64_t>( n, (c + 1) * (uint64_t)CHUNK );
			Ref* refs = buf[0].get();
			for (uint32_t i = (uint32_t)c * CHUNK; i < end; i++)
			{
				Ref& r = refs[i];
				for (int a = 0; a < 3; a++) r.lo[a] = primBoxes[i].lo[a], r.hi[a] = primBoxes[i].hi[a];
				r.id = i, r.pad = 0;
			}
			boundsOf( refs + c * CHUNK, end - (uint32_t)c * CHUNK, box, cbox );
			part[2 * c] = box, part[2 * c + 1] = cbox;
		}, 2 );
		Box4 rootBox, rootCbox;
		rootBox.reset(), rootCbox.reset();
		for (uint32_t c = 0; c < rootChunks; c++) rootBox.grow( part[2 * c] ), rootCbox.grow( part[2 * c + 1] );
		const Work root = work( 0, 0, n, 0, rootBox, rootCbox );
		// ---- top of the tree, a level at a time: ranges of >= big references are binned and split by all threads together
		const uint32_t big = threads > 1 ? std::max<uint32_t>( 4 * CHUNK, n / (8 * (uint32_t)threads) ) : 0xffffffffu;
		std::vector<Work> level{ root }, nextLevel, tasks;
		std::vector<uint32_t> topNodes; // in creation order: parents before children
		struct Chunk { uint32_t work, begin, count, left, right; }; // left / right: where this chunk's halves start in the other buffer
		std::vector<Chunk> chunks;
		std::vector<Bins> chunkBins;
		std::vector<HalfBoxes> chunkBoxes;
		std::vector<BinMap> maps;
		std::vector<Split> splits;
		std::unique_ptr<Bins> total( new Bins );
		while (!level.empty())
		{
			std::vector<Work> wide;
			for (const Work& w : level) if (w.count >= big) wide.push_back( w ); else tasks.push_back( w );
			if (wide.empty()) break;
			chunks.clear(), maps.clear(), splits.clear(), nextLevel.clear();
			for (size_t i = 0; i < wide.size(); i++)
			{
				initNode( wide[i] ), topNodes.push_back( wide[i].node );
				maps.emplace_back( wide[i].cbox, binsFor( wide[i].count ) );
				for (uint32_t b = 0; b < wide[i].count; b += CHUNK) chunks.push_back( { (uint32_t)i, wide[i].first + b, std::min( CHUNK, wide[i].count - b ), 0, 0 } );
			}
			if (chunkBins.size() < chunks.size()) chunkBins.resize( chunks.size() ), chunkBoxes.resize( chunks.size() );
			ParallelFor( threads, chunks.size(), [&]( size_t c ) {
				const Chunk& ch = chunks[c];
				chunkBins[c].clear();
				binRange( buf[wide[ch.work].src].get() + ch.begin, ch.count, maps[ch.work], chunkBins[c] );
			}, 2 );
			for (size_t i = 0, c = 0; i < wide.size(); i++)
			{
				total->clear();
				const size_t c0 = c;
				for (; c < chunks.size() && chunks[c].work == i; c++) total->add( chunkBins[c] );
				splits.push_back( chooseSplit( *total, wide[i].cbox ) );
				const Split& s = splits.back();
				if (s.axis < 0) continue;
				uint32_t left = wide[i].first, right = wide[i].first + s.leftCount;
				for (size_t k = c0; k < c; k++)
				{
					uint32_t lc = 0;
					for (int b = 0; b <= s.bin; b++) lc += chunkBins[k].count[s.axis][b];
					chunks[k].left = left, chunks[k].right = right;
					left += lc, right += chunks[k].count - lc;
				}
			}
			ParallelFor( threads, chunks.size(), [&]( size_t c ) {
				const Chunk& ch = chunks[c];
				chunkBoxes[c].reset();
				if (splits[ch.work].axis < 0) return;
				Ref* dst = buf[wide[ch.work].src ^ 1].get();
				splitRange( buf[wide[ch.work].src].get() + ch.begin, ch.count, maps[ch.work], splits[ch.work], dst + ch.left, dst + ch.right, chunkBoxes[c] );
			}, 2 );
			for (size_t i = 0, c = 0; i < wide.size(); i++)
			{
				HalfBoxes hb;
				hb.reset();
				for (; c < chunks.size() && chunks[c].work == i; c++)
					for (int s = 0; s < 2; s++) hb.box[s].grow( chunkBoxes[c].box[s] ), hb.cbox[s].grow( chunkBoxes[c].cbox[s] );
				Work child[2];
				if (splits[i].axis >= 0) children( wide[i], splits[i], hb, child, 0 );
				else if (!degenerate( wide[i], child, 0 )) continue;
				nextLevel.push_back( child[0] ), nextLevel.push_back( child[1] );
			}
			level.swap( nextLevel );
		}
		// ---- the subtrees below, largest first
		std::stable_sort( tasks.begin(), tasks.end(), []( const Work& a, const Work& b ) { return a.count > b.count; } );
		ParallelFor( threads, tasks.size(), [&]( size_t t ) { buildTask( tasks[t] ); } );
		for (size_t i = topNodes.size(); i-- > 0;) plan.node( nodes.get(), topNodes[i], maxLeaf );
	}
};

} // namespace

void BuildWideBvh( const Aabb* primBoxes, uint32_t n, int maxLeafPrims, int threads, WideBvh& out )
{
	out.nodes.clear(), out.levelStart.clear(), out.primOrder.clear();
	out.bounds.reset();
	if (n == 0)
	{
		// an empty tree is one node without children: every ray misses
		WideNode wn;
		memset( &wn, 0, sizeof( wn ) );
		wn.e[0] = wn.e[1] = wn.e[2] = 27;
		out.nodes.push_back( wn );
		out.levelStart = { 0, 1 };
		return;
	}
	// B200_BUILD_TIMING=1: phase times of large builds on stderr
	static const bool timing = getenv( "B200_BUILD_TIMING" ) != 0;
	auto now = [] { return std::chrono::steady_clock::now(); };
	auto seconds = []( std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b ) { return std::chrono::duration<double>( b - a ).count(); };
	const auto t0 = now();
	if (threads > 1 && n < 16384) threads = 1; // the top level and small meshes: not worth a thread
	Builder2 b2;
	b2.maxLeaf = maxLeafPrims, b2.threads = threads;
	b2.build( primBoxes, n );
	const auto t1 = now();
	const Node2* N2 = b2.nodes.get();
	const Ref* refs = b2.buf[0].get();
	const CollapsePlan& plan = b2.plan;
	out.bounds = N2[0].b;

	// Breadth-first emission, a level at a time.  Per level: (1) every wide node of the level picks its children (the forest the
	// SAH-optimal collapse chose for it) and their slots, (2) a running sum over the level places the children's nodes and the leaves'
	// primitives, (3) every node is encoded and its primitives are listed.  (1) and (3) run on all threads; the result is that of the
	// one-node-at-a-time loop.
	struct Item {
		uint32_t wide, n2;          // wide node index, BVH2 node it stands for
		uint32_t child[8];          // BVH2 node per slot, ~0u = empty
		uint32_t childBase, primBase, innerCount, primCount;
	};
	std::vector<Item> cur( 1 ), next;
	cur[0].wide = 0, cur[0].n2 = 0;
	out.nodes.resize( 1 );
	out.primOrder.resize( n );
	out.levelStart.push_back( 0 );
	uint32_t nodeEnd = 1, primEnd = 0;
	auto isInner = [&]( uint32_t m ) { return N2[m].count == 0 && !plan.asLeaf[m]; };
	constexpr size_t GRAIN = 256;
	while (!cur.empty())
	{
		const int levelThreads = cur.size() >= 4 * GRAIN ? threads : 1;
		const size_t grains = (cur.size() + GRAIN - 1) / GRAIN;
		ParallelFor( levelThreads, grains, [&]( size_t g ) {
			for (size_t idx = g * GRAIN; idx < std::min( cur.size(), (g + 1) * GRAIN ); idx++)
			{
				Item& it = cur[idx];
				uint32_t child[8];
				int cn = 0;
				const Node2& root = N2[it.n2];
				if (root.count > 0 || plan.asLeaf[it.n2]) child[cn++] = it.n2; // the whole tree is one leaf
				else
				{
					const int k = plan.split8[it.n2];
					plan.expand( N2, root.left, k, child, cn );
					plan.expand( N2, root.right, 8 - k, child, cn );
				}
				Aabb cb[8];
				for (int i = 0; i < cn; i++) cb[i] = N2[child[i]].b;
				int slotOf[8];
				AssignSlots( root.b, cb, cn, slotOf );
				for (int s = 0; s < 8; s++) it.child[s] = ~0u;
				it.innerCount = it.primCount = 0;
				for (int i = 0; i < cn; i++)
				{
					it.child[slotOf[i]] = child[i];
					if (isInner( child[i] )) it.innerCount++; else it.primCount += N2[child[i]].span;
				}
			}
		} );
		size_t nextCount = 0;
		for (Item& it : cur) it.childBase = nodeEnd, it.primBase = primEnd, nodeEnd += it.innerCount, primEnd += it.primCount, nextCount += it.innerCount;
		out.nodes.resize( nodeEnd );
		next.resize( nextCount );
		const uint32_t nextStart = nodeEnd - (uint32_t)nextCount;
		ParallelFor( levelThreads, grains, [&]( size_t g ) {
			for (size_t idx = g * GRAIN; idx < std::min( cur.size(), (g + 1) * GRAIN ); idx++)
			{
				const Item& it = cur[idx];
				ChildDesc desc[8];
				const ChildDesc* slot[8];
				uint32_t innerCount = 0, primOffset = 0;
				for (int s = 0; s < 8; s++)
				{
					slot[s] = 0;
					if (it.child[s] == ~0u) continue;
					const Node2& c = N2[it.child[s]];
					const bool inner = isInner( it.child[s] );
					desc[s].b = c.b, desc[s].inner = inner, desc[s].primCount = inner ? 0 : c.span, desc[s].primOffset = primOffset;
					slot[s] = &desc[s];
					if (inner)
					{
						Item& nx = next[it.childBase + innerCount - nextStart];
						nx.wide = it.childBase + innerCount, nx.n2 = it.child[s];
						innerCount++;
					}
					else
					{
						for (uint32_t k = 0; k < c.span; k++) out.primOrder[it.primBase + primOffset + k] = refs[c.first + k].id;
						primOffset += c.span;
					}
				}
				WideNode& wn = out.nodes[it.wide];
				memset( &wn, 0, sizeof( wn ) );
				EncodeNode( wn, N2[it.n2].b, slot );
				wn.childBase = it.childBase, wn.primBase = it.primBase;
			}
		} );
		out.levelStart.push_back( nextStart );
		cur.swap( next );
	}
	// levelStart currently holds starts of each level plus the start of the (empty) level after the last
	out.levelStart.back() = (uint32_t)out.nodes.size();
	if (timing && n >= 65536)
		fprintf( stderr, "[b200 build] %u prims, %d threads: binary tree + collapse plan %.3f s, emission %.3f s\n", n, threads, seconds( t0, t1 ), seconds( t1, now() ) );
}


} // namespace b200

// ---- host refit (reference implementation of the GPU refit kernels in refit.cu) ---------------------
namespace b200 {

void RefitWideBvhHost( WideBvh& bvh, const Aabb* primBoxes )
{
	std::vector<Aabb> nodeBox( bvh.nodes.size() );
	for (int level = (int)bvh.levelStart.size() - 2; level >= 0; level--)
		for (uint32_t ni = bvh.levelStart[level]; ni < bvh.levelStart[level + 1]; ni++)
		{
			WideNode& wn = bvh.nodes[ni];
			Aabb cb[8], box;
			box.reset();
			uint32_t inner = 0;
			for (int s = 0; s < 8; s++)
			{
				cb[s].reset();
				uint8_t m = wn.meta[s];
				if (!m) continue;
				if (wn.imask & (1u << s)) cb[s] = nodeBox[wn.childBase + inner++];
				else
				{
					uint32_t cnt = __builtin_popcount( m >> 5 ), off = m & 31;
					for (uint32_t k = 0; k < cnt; k++) cb[s].grow( primBoxes[bvh.primOrder[wn.primBase + off + k]] );
				}
				box.grow( cb[s] );
			}
			nodeBox[ni] = box;
			QuantiseFrame( box, wn.p, wn.e );
			for (int s = 0; s < 8; s++)
			{
				if (!wn.meta[s]) continue;
				uint8_t qlo[3], qhi[3];
				QuantiseChild( wn.p, wn.e, cb[s], qlo, qhi );
				wn.qlox[s] = qlo[0], wn.qloy[s] = qlo[1], wn.qloz[s] = qlo[2];
				wn.qhix[s] = qhi[0], wn.qhiy[s] = qhi[1], wn.qhiz[s] = qhi[2];
			}
		}
	if (!nodeBox.empty() && !bvh.primOrder.empty()) bvh.bounds = nodeBox[0];
}

} // namespace b200
