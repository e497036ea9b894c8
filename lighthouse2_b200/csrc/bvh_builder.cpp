// bvh_builder.cpp -- host-side converter: primitive boxes -> binned-SAH BVH2 -> compressed 8-wide BVH.
//
// This is RenderCore_B200's own builder (it replaces BVH2::Rebuild, lib/RenderCore_Bart/bvh.cpp:57-256,
// and optixAccelBuild, lib/rendercore_optix7/core_mesh.cpp:94-130).  The tree it produces is NOT the
// reference's tree and does not need to be: the reference BLAS traversal never culls by distance
// (bvh.cpp:281-298), so any conservative hierarchy yields the same closest hit.
#include "wide_bvh.h"
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstring>
#include <thread>

namespace b200 {

float g_sahPrimCost = 0.4f;

namespace {

struct Ref { Aabb b; float c[3]; uint32_t id; }; // primitive reference, partitioned in place

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

struct Builder2 {
	std::vector<Ref> refs;
	std::vector<Node2> nodes;
	std::atomic<uint32_t> nodeCount{ 0 };
	std::atomic<int> spareThreads{ 0 };
	int maxLeaf = 3;

	uint32_t alloc2() { return nodeCount.fetch_add( 2 ); }

	void build( uint32_t nodeIdx, uint32_t first, uint32_t count )
	{
		Node2& node = nodes[nodeIdx];
		Aabb box, cbox;
		box.reset(), cbox.reset();
		for (uint32_t i = first; i < first + count; i++) box.grow( refs[i].b ), cbox.grow( refs[i].c );
		node.b = box;
		node.first = first, node.count = count, node.span = count, node.left = node.right = 0;
		if (count == 1) return;
		// ---- binned SAH over the three axes of the centroid box
		float bestCost = 3.0e38f;
		const int nb = count >= 256 ? BINS : (BINS < 16 ? BINS : 16); // the fine bins pay at the top of the tree, where the ranges are long
		int bestAxis = -1, bestSplit = 0;
		for (int a = 0; a < 3; a++)
		{
			float ext = cbox.hi[a] - cbox.lo[a];
			if (!(ext > 0)) continue;
			Aabb binBox[BINS];
			uint32_t binCount[BINS];
			for (int k = 0; k < nb; k++) binBox[k].reset(), binCount[k] = 0;
			float scale = (float)nb / ext;
			for (uint32_t i = first; i < first + count; i++)
			{
				int k = (int)((refs[i].c[a] - cbox.lo[a]) * scale);
				k = k < 0 ? 0 : (k > nb - 1 ? nb - 1 : k);
				binBox[k].grow( refs[i].b ), binCount[k]++;
			}
			float rightArea[BINS];
			uint32_t rightCount[BINS];
			Aabb acc;
			acc.reset();
			uint32_t n = 0;
			for (int k = nb - 1; k > 0; k--)
			{
				if (binCount[k]) acc.grow( binBox[k] );
				n += binCount[k];
				rightArea[k] = n ? acc.halfArea() : 0, rightCount[k] = n;
			}
			acc.reset(), n = 0;
			for (int k = 0; k < nb - 1; k++)
			{
				if (binCount[k]) acc.grow( binBox[k] );
				n += binCount[k];
				if (n == 0 || rightCount[k + 1] == 0) continue;
				float cost = acc.halfArea() * n + rightArea[k + 1] * rightCount[k + 1];
				if (cost < bestCost) bestCost = cost, bestAxis = a, bestSplit = k;
			}
		}
		uint32_t mid;
		if (bestAxis >= 0)
		{
			float splitCost = COST_TRAVERSE + COST_PRIM * bestCost / std::max( box.halfArea(), 1e-30f );
			if (!B200_DP_LEAVES && (int)count <= maxLeaf && (float)count * COST_PRIM <= splitCost) return; // cheaper as a leaf
			float scale = (float)nb / (cbox.hi[bestAxis] - cbox.lo[bestAxis]);
			float lo = cbox.lo[bestAxis];
			int a = bestAxis, s = bestSplit;
			auto it = std::partition( refs.begin() + first, refs.begin() + first + count, [=]( const Ref& r ) {
				int k = (int)((r.c[a] - lo) * scale);
				k = k < 0 ? 0 : (k > nb - 1 ? nb - 1 : k);
				return k <= s;
			} );
			mid = (uint32_t)(it - refs.begin());
			if (mid == first || mid == first + count) mid = first + count / 2; // numerically degenerate: median by index
		}
		else
		{
			if ((int)count <= maxLeaf) return; // identical centroids and small enough
			mid = first + count / 2;
		}
		uint32_t l = alloc2();
		node.left = l, node.right = l + 1, node.count = 0;
		uint32_t lc = mid - first, rc = count - lc;
		// fork the larger subtrees while spare hardware threads exist
		if (lc > 32768 && rc > 32768 && spareThreads.fetch_sub( 1 ) > 0)
		{
			std::thread t( [=]() { build( l, first, lc ); spareThreads.fetch_add( 1 ); } );
			build( l + 1, mid, rc );
			t.join();
		}
		else
		{
			if (lc > 32768 && rc > 32768) spareThreads.fetch_add( 1 ); // undo the failed reservation
			build( l, first, lc );
			build( l + 1, mid, rc );
		}
	}
};

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
// COST_PRIM.  BVH2 children always have larger indices than their parent, so one descending sweep suffices.
struct CollapsePlan {
	std::vector<float> cost;      // [n*7 + (i-1)]
	std::vector<uint8_t> choice;  // [n*7 + (i-1)], i >= 2: 0 = "use i-1 roots", k>0 = left gets k roots, right i-k
	std::vector<uint8_t> split8;  // [n]: left's share of the 8 child slots when n becomes a wide node
	std::vector<uint8_t> asLeaf;  // [n]: an inner BVH2 node of <= maxLeaf primitives that is cheaper as ONE leaf (B200_DP_LEAVES)

	void compute( const std::vector<Node2>& N2, uint32_t used, int maxLeaf )
	{
		cost.assign( (size_t)used * 7, 0.f ), choice.assign( (size_t)used * 7, 0 ), split8.assign( used, 0 ), asLeaf.assign( used, 0 );
		for (int64_t n = (int64_t)used - 1; n >= 0; n--)
		{
			const Node2& nd = N2[n];
			float A = nd.b.halfArea();
			float* c = &cost[(size_t)n * 7];
			if (nd.count > 0) { for (int i = 0; i < 7; i++) c[i] = A * nd.count * COST_PRIM; continue; }
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
			for (int i = 2; i <= 7; i++)
			{
				int k = 1;
				float d = distribute( i, k );
				if (d < c[i - 2]) c[i - 1] = d, choice[(size_t)n * 7 + i - 1] = (uint8_t)k;
				else c[i - 1] = c[i - 2], choice[(size_t)n * 7 + i - 1] = 0;
			}
		}
	}
	// the roots of the forest chosen for BVH2 node m when it may use up to i child slots
	void expand( const std::vector<Node2>& N2, uint32_t m, int i, uint32_t* out, int& cn ) const
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
	Builder2 b2;
	b2.maxLeaf = maxLeafPrims;
	b2.refs.resize( n );
	for (uint32_t i = 0; i < n; i++)
	{
		Ref& r = b2.refs[i];
		r.b = primBoxes[i], r.id = i;
		for (int a = 0; a < 3; a++) r.c[a] = 0.5f * (r.b.lo[a] + r.b.hi[a]);
	}
	b2.nodes.resize( 2 * (size_t)n + 2 );
	b2.nodeCount = 1;
	b2.spareThreads = threads > 1 ? threads - 1 : 0;
	b2.build( 0, 0, n );
	const std::vector<Node2>& N2 = b2.nodes;
	out.bounds = N2[0].b;

	CollapsePlan plan;
	plan.compute( N2, b2.nodeCount.load(), maxLeafPrims );

	// breadth-first emission: queue entries = (wide node index, BVH2 node index)
	struct Item { uint32_t wide, n2; };
	std::vector<Item> cur, next;
	out.nodes.reserve( n / 3 + 16 );
	out.primOrder.reserve( n );
	out.nodes.emplace_back();
	cur.push_back( { 0, 0 } );
	out.levelStart.push_back( 0 );
	while (!cur.empty())
	{
		next.clear();
		for (const Item& it : cur)
		{
			// children of this wide node = the forest the SAH-optimal collapse (below) chose for it
			uint32_t child[8];
			int cn = 0;
			const Node2& root = N2[it.n2];
			if (root.count > 0 || plan.asLeaf[it.n2]) child[cn++] = it.n2; // the whole tree is one leaf
			else
			{
				int k = plan.split8[it.n2];
				plan.expand( N2, root.left, k, child, cn );
				plan.expand( N2, root.right, 8 - k, child, cn );
			}
			Aabb cb[8];
			for (int i = 0; i < cn; i++) cb[i] = N2[child[i]].b;
			int slotOf[8];
			AssignSlots( root.b, cb, cn, slotOf );
			int childInSlot[8];
			for (int s = 0; s < 8; s++) childInSlot[s] = -1;
			for (int i = 0; i < cn; i++) childInSlot[slotOf[i]] = i;
			ChildDesc desc[8];
			const ChildDesc* slot[8];
			uint32_t childBase = (uint32_t)out.nodes.size(), primBase = (uint32_t)out.primOrder.size();
			uint32_t innerCount = 0, primOffset = 0;
			for (int s = 0; s < 8; s++)
			{
				slot[s] = 0;
				int i = childInSlot[s];
				if (i < 0) continue;
				const Node2& c = N2[child[i]];
				const bool inner = c.count == 0 && !plan.asLeaf[child[i]];
				desc[s].b = c.b, desc[s].inner = inner, desc[s].primCount = inner ? 0 : c.span, desc[s].primOffset = primOffset;
				slot[s] = &desc[s];
				if (inner)
				{
					next.push_back( { childBase + innerCount, child[i] } );
					innerCount++;
				}
				else
				{
					for (uint32_t k = 0; k < c.span; k++) out.primOrder.push_back( b2.refs[c.first + k].id );
					primOffset += c.span;
				}
			}
			out.nodes.resize( out.nodes.size() + innerCount );
			WideNode& wn = out.nodes[it.wide];
			memset( &wn, 0, sizeof( wn ) );
			EncodeNode( wn, root.b, slot );
			wn.childBase = childBase, wn.primBase = primBase;
		}
		out.levelStart.push_back( (uint32_t)(out.nodes.size() - next.size()) );
		cur.swap( next );
	}
	// levelStart currently holds starts of each level plus the start of the (empty) level after the last
	out.levelStart.back() = (uint32_t)out.nodes.size();
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
