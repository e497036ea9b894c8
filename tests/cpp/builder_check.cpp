// tests/cpp/builder_check.cpp -- host-only check of the BVH builder (lighthouse2_b200/csrc/bvh_builder.cpp), no CUDA needed.
// Compiled and run by tests/test_builder_host.py:  builder_check <case> <n> <threads>  prints one line of key=value pairs.
//
// For the tree of one input it verifies what the traversal kernels rely on (wide_bvh.h):
//   * every primitive is listed exactly once in leaf order; leaves hold 1..3 primitives inside their node's window of 24;
//   * inner children sit contiguously at childBase in ascending slot order, levels are contiguous (breadth-first) and every node but
//     the root has exactly one parent;
//   * conservativeness: the decoded 8-bit planes of every child slot contain the boxes of all primitives below it;
//   * the tree does not depend on the number of threads that built it (the caller compares the hashes of two runs).
#include "wide_bvh.h"
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
using namespace b200;

static std::vector<Aabb> makeCase( const std::string& name, uint32_t n )
{
	std::mt19937 rng( 12345 );
	std::uniform_real_distribution<float> U( 0, 1 );
	std::vector<Aabb> boxes( n );
	auto triBox = [&]( float cx, float cy, float cz, float size ) {
		Aabb b;
		b.reset();
		for (int k = 0; k < 3; k++)
		{
			float p[3] = { cx + (U( rng ) - 0.5f) * size, cy + (U( rng ) - 0.5f) * size, cz + (U( rng ) - 0.5f) * size };
			b.grow( p );
		}
		return b;
	};
	for (uint32_t i = 0; i < n; i++)
	{
		if (name == "soup") boxes[i] = triBox( U( rng ) * 10, U( rng ) * 10, U( rng ) * 10, 0.05f );
		else if (name == "grid")
		{
			// a tessellated plane: rows of equal centroid coordinates, flat on one axis
			const uint32_t w = (uint32_t)std::ceil( std::sqrt( (double)n ) );
			const float x = (float)(i % w), z = (float)(i / w);
			Aabb b;
			b.lo[0] = x, b.hi[0] = x + 1, b.lo[1] = b.hi[1] = 0.5f, b.lo[2] = z, b.hi[2] = z + 1;
			boxes[i] = b;
		}
		else if (name == "same")
		{
			// every primitive in the same place: no plane separates the centroids anywhere
			Aabb b;
			for (int a = 0; a < 3; a++) b.lo[a] = 1.0f, b.hi[a] = 2.0f;
			boxes[i] = b;
		}
		else if (name == "clusters")
		{
			// a few far-apart clumps with many coincident primitives in each, sizes spanning 2^-20 .. 2^10
			const int c = (int)(rng() % 7);
			const float size = std::ldexp( 1.0f, (int)(rng() % 31) - 20 );
			if (rng() % 3 == 0) boxes[i] = triBox( c * 1000.0f, -c * 500.0f, 3.0f, 0 );
			else boxes[i] = triBox( c * 1000.0f + U( rng ), -c * 500.0f + U( rng ), 3.0f + U( rng ), size );
		}
		else { fprintf( stderr, "unknown case %s\n", name.c_str() ); exit( 2 ); }
	}
	return boxes;
}

struct Checker {
	const WideBvh& bvh;
	const std::vector<Aabb>& boxes;
	std::vector<uint32_t> parents, seen;
	uint64_t slots = 0, leaves = 0;
	std::string error;
	Checker( const WideBvh& b, const std::vector<Aabb>& bx ) : bvh( b ), boxes( bx ), parents( b.nodes.size(), 0 ), seen( bx.size(), 0 ) {}
	bool fail( const char* what, uint32_t node ) { if (error.empty()) error = std::string( what ) + " at node " + std::to_string( node ); return false; }
	// returns the box of everything below `node`
	bool walk( uint32_t node, int level, Aabb& out )
	{
		if (node >= bvh.nodes.size()) return fail( "child index out of range", node );
		if (level + 1 >= (int)bvh.levelStart.size() || node < bvh.levelStart[level] || node >= bvh.levelStart[level + 1]) return fail( "node outside its level", node );
		const WideNode& wn = bvh.nodes[node];
		out.reset();
		uint32_t inner = 0;
		for (int s = 0; s < 8; s++)
		{
			const uint8_t meta = wn.meta[s];
			const bool isInner = (wn.imask >> s) & 1;
			if (meta == 0) { if (isInner) return fail( "inner bit on an empty slot", node ); continue; }
			slots++;
			Aabb below;
			below.reset();
			if (isInner)
			{
				if (meta != (0x20 | (24 + s))) return fail( "inner meta byte", node );
				const uint32_t child = wn.childBase + inner++;
				if (child < bvh.nodes.size()) parents[child]++;
				if (!walk( child, level + 1, below )) return false;
			}
			else
			{
				const int unary = meta >> 5, offset = meta & 31;
				const int count = unary == 1 ? 1 : unary == 3 ? 2 : unary == 7 ? 3 : -1;
				if (count < 0 || offset + count > 24) return fail( "leaf meta byte", node );
				leaves++;
				for (int k = 0; k < count; k++)
				{
					const uint64_t at = (uint64_t)wn.primBase + offset + k;
					if (at >= bvh.primOrder.size() || bvh.primOrder[at] >= boxes.size()) return fail( "primitive slot out of range", node );
					seen[bvh.primOrder[at]]++;
					below.grow( boxes[bvh.primOrder[at]] );
				}
			}
			// decoded planes of the slot (what the kernels intersect) must contain everything below it
			const uint8_t qlo[3] = { wn.qlox[s], wn.qloy[s], wn.qloz[s] }, qhi[3] = { wn.qhix[s], wn.qhiy[s], wn.qhiz[s] };
			for (int a = 0; a < 3; a++)
			{
				const double cell = std::ldexp( 1.0, (int)wn.e[a] - 127 );
				const double lo = (double)wn.p[a] + qlo[a] * cell, hi = (double)wn.p[a] + qhi[a] * cell;
				if (!(lo <= below.lo[a] && hi >= below.hi[a])) return fail( "child planes do not contain the geometry below", node );
			}
			out.grow( below );
		}
		return true;
	}
};

int main( int argc, char** argv )
{
	if (argc < 4) { fprintf( stderr, "usage: builder_check <soup|grid|same|clusters> <n> <threads> [maxLeaf]\n" ); return 2; }
	const std::string name = argv[1];
	const uint32_t n = (uint32_t)atoll( argv[2] );
	const int threads = atoi( argv[3] ), maxLeaf = argc > 4 ? atoi( argv[4] ) : 3;
	const std::vector<Aabb> boxes = makeCase( name, n );
	WideBvh bvh;
	BuildWideBvh( boxes.data(), n, maxLeaf, threads, bvh );
	Checker check( bvh, boxes );
	Aabb all;
	bool ok = bvh.levelStart.size() >= 2 && bvh.levelStart[0] == 0 && bvh.levelStart[1] == 1 && bvh.levelStart.back() == bvh.nodes.size();
	if (!ok) check.error = "level table";
	if (ok && n) ok = check.walk( 0, 0, all );
	if (ok && bvh.primOrder.size() != n) ok = false, check.error = "primitive order has the wrong length";
	for (uint32_t i = 0; ok && i < n; i++) if (check.seen[i] != 1) ok = false, check.error = "primitive " + std::to_string( i ) + " listed " + std::to_string( check.seen[i] ) + " times";
	for (size_t i = 1; ok && i < bvh.nodes.size(); i++) if (check.parents[i] != 1) ok = false, check.error = "node " + std::to_string( i ) + " has " + std::to_string( check.parents[i] ) + " parents";
	for (int a = 0; ok && n && a < 3; a++) if (bvh.bounds.lo[a] != all.lo[a] || bvh.bounds.hi[a] != all.hi[a]) ok = false, check.error = "bounds";
	uint64_t h = 1469598103934665603ull;
	const unsigned char* p = (const unsigned char*)bvh.nodes.data();
	for (size_t i = 0; i < bvh.nodes.size() * sizeof( WideNode ); i++) h = (h ^ p[i]) * 1099511628211ull;
	for (uint32_t v : bvh.primOrder) h = (h ^ v) * 1099511628211ull;
	for (uint32_t v : bvh.levelStart) h = (h ^ v) * 1099511628211ull;
	printf( "ok=%d nodes=%zu levels=%zu slots=%llu leaves=%llu hash=%016llx error=%s\n", ok ? 1 : 0, bvh.nodes.size(), bvh.levelStart.size() - 1,
		(unsigned long long)check.slots, (unsigned long long)check.leaves, (unsigned long long)h, check.error.empty() ? "none" : check.error.c_str() );
	return ok ? 0 : 1;
}
