// bvh_cache.cpp -- on-disk cache of host-built wide BVHs.  The reference tracks changes of every scene object
// with a CRC64 (TRACKCHANGES, lib/platform/system.h); here the key is a 64-bit hash of the mesh's vertex data and of the builder
// parameters, the value is what BuildWideBvh produced: nodes, level table, primitive order, bounds.  One file per mesh in the
// directory named by the environment variable B200_BVH_CACHE; a file that fails any check is ignored and rebuilt.
#include "wide_bvh.h"
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

namespace b200 {

namespace {
constexpr uint32_t MAGIC = 0x42573242u /* "B2WB" */, VERSION = 4; // 3: 64 SAH bins, leaves chosen by the collapse; 4: node frames with headroom (QuantiseFrame)
struct Header {
	uint32_t magic, version;
	uint64_t key;
	uint32_t triCount, nodeCount, levelCount, maxLeafPrims;
	float sahPrimCost, lo[3], hi[3];
	uint64_t payloadHash;
};
inline uint64_t mixWord( uint64_t h, const uint64_t w )
{
	h ^= w, h *= 0x9E3779B97F4A7C15ull, h ^= h >> 29;
	return h;
}
}

// four independent lanes over 8-byte words (the loop carries no dependency between lanes), folded at the end
uint64_t HashBytes( const void* data, size_t bytes, uint64_t seed )
{
	const unsigned char* p = (const unsigned char*)data;
	uint64_t h[4] = { seed ^ 0x243F6A8885A308D3ull, seed ^ 0x13198A2E03707344ull, seed ^ 0xA4093822299F31D0ull, seed ^ 0x082EFA98EC4E6C89ull };
	size_t i = 0;
	for (; i + 32 <= bytes; i += 32)
	{
		uint64_t w[4];
		memcpy( w, p + i, 32 );
		h[0] = mixWord( h[0], w[0] ), h[1] = mixWord( h[1], w[1] ), h[2] = mixWord( h[2], w[2] ), h[3] = mixWord( h[3], w[3] );
	}
	uint64_t tail[4] = { 0, 0, 0, 0 };
	memcpy( tail, p + i, bytes - i );
	for (int k = 0; k < 4; k++) h[k] = mixWord( h[k], tail[k] );
	uint64_t r = mixWord( seed, (uint64_t)bytes );
	for (int k = 0; k < 4; k++) r = mixWord( r, h[k] );
	return r;
}

static std::string cachePath( const std::string& dir, uint64_t key, uint32_t T )
{
	char name[64];
	snprintf( name, sizeof( name ), "/widebvh_%016llx_%u.b2wb", (unsigned long long)key, T );
	return dir + name;
}

bool LoadWideBvh( const std::string& dir, uint64_t key, uint32_t T, int maxLeafPrims, WideBvh& out )
{
	FILE* f = fopen( cachePath( dir, key, T ).c_str(), "rb" );
	if (!f) return false;
	Header h;
	bool ok = fread( &h, sizeof( h ), 1, f ) == 1 && h.magic == MAGIC && h.version == VERSION && h.key == key && h.triCount == T &&
		h.maxLeafPrims == (uint32_t)maxLeafPrims && h.sahPrimCost == g_sahPrimCost && h.nodeCount >= 1 && h.nodeCount <= T + 64 && h.levelCount >= 1 && h.levelCount <= 128;
	if (ok)
	{
		out.nodes.resize( h.nodeCount ), out.levelStart.resize( h.levelCount + 1 ), out.primOrder.resize( T );
		ok = fread( out.levelStart.data(), 4, out.levelStart.size(), f ) == out.levelStart.size() && fread( out.primOrder.data(), 4, T, f ) == T &&
			 fread( out.nodes.data(), sizeof( WideNode ), h.nodeCount, f ) == h.nodeCount;
	}
	fclose( f );
	if (!ok) return false;
	// the payload must be what was written, and must not be able to send the kernels out of bounds
	uint64_t ph = HashBytes( out.levelStart.data(), out.levelStart.size() * 4, 1 );
	ph = HashBytes( out.primOrder.data(), (size_t)T * 4, ph ), ph = HashBytes( out.nodes.data(), (size_t)h.nodeCount * sizeof( WideNode ), ph );
	if (ph != h.payloadHash) return false;
	if (out.levelStart.front() != 0 || out.levelStart.back() != h.nodeCount) return false;
	for (size_t i = 1; i < out.levelStart.size(); i++) if (out.levelStart[i] < out.levelStart[i - 1]) return false;
	std::vector<bool> seen( T, false );
	for (uint32_t i = 0; i < T; i++)
	{
		const uint32_t p = out.primOrder[i];
		if (p >= T || seen[p]) return false;
		seen[p] = true;
	}
	// per node: the inner children (one per imask bit, contiguous at childBase) must be nodes of the NEXT level, every leaf slot's
	// triangle run must end inside the triangle array, slot bytes must be well formed; and the tree must fit the traversal stack
	if (h.levelCount > MAX_TREE_LEVELS) return false;
	for (uint32_t level = 0; level + 1 < (uint32_t)out.levelStart.size(); level++)
		for (uint32_t ni = out.levelStart[level]; ni < out.levelStart[level + 1]; ni++)
		{
			const WideNode& n = out.nodes[ni];
			uint32_t inner = 0, primEnd = 0;
			for (int s = 0; s < 8; s++)
			{
				const uint32_t meta = n.meta[s];
				const bool isInner = (n.imask >> s) & 1u;
				if (!meta) { if (isInner) return false; continue; }
				if (isInner) { if (meta != (0x20u | (24u + (uint32_t)s))) return false; inner++; continue; }
				const uint32_t unary = meta >> 5, off = meta & 31u;
				if (unary != 1 && unary != 3 && unary != 7) return false;
				const uint32_t cnt = unary == 1 ? 1 : unary == 3 ? 2 : 3;
				if ((int)cnt > maxLeafPrims || off + cnt > 24) return false;
				if (off + cnt > primEnd) primEnd = off + cnt;
			}
			if (primEnd && ((uint64_t)n.primBase + primEnd > T)) return false;
			if (inner)
			{
				if (level + 2 >= (uint32_t)out.levelStart.size()) return false; // the last level has no children
				if (n.childBase < out.levelStart[level + 1] || (uint64_t)n.childBase + inner > out.levelStart[level + 2]) return false;
			}
		}
	for (int a = 0; a < 3; a++) out.bounds.lo[a] = h.lo[a], out.bounds.hi[a] = h.hi[a];
	return true;
}

bool StoreWideBvh( const std::string& dir, uint64_t key, uint32_t T, int maxLeafPrims, const WideBvh& bvh )
{
	Header h;
	memset( &h, 0, sizeof( h ) );
	h.magic = MAGIC, h.version = VERSION, h.key = key, h.triCount = T, h.nodeCount = (uint32_t)bvh.nodes.size();
	h.levelCount = (uint32_t)bvh.levelStart.size() - 1, h.maxLeafPrims = (uint32_t)maxLeafPrims, h.sahPrimCost = g_sahPrimCost;
	for (int a = 0; a < 3; a++) h.lo[a] = bvh.bounds.lo[a], h.hi[a] = bvh.bounds.hi[a];
	uint64_t ph = HashBytes( bvh.levelStart.data(), bvh.levelStart.size() * 4, 1 );
	ph = HashBytes( bvh.primOrder.data(), (size_t)T * 4, ph ), ph = HashBytes( bvh.nodes.data(), bvh.nodes.size() * sizeof( WideNode ), ph );
	h.payloadHash = ph;
	const std::string path = cachePath( dir, key, T ), tmp = path + ".tmp";
	FILE* f = fopen( tmp.c_str(), "wb" );
	if (!f) return false;
	const bool ok = fwrite( &h, sizeof( h ), 1, f ) == 1 && fwrite( bvh.levelStart.data(), 4, bvh.levelStart.size(), f ) == bvh.levelStart.size() &&
		fwrite( bvh.primOrder.data(), 4, T, f ) == T && fwrite( bvh.nodes.data(), sizeof( WideNode ), bvh.nodes.size(), f ) == bvh.nodes.size();
	if (fclose( f ) != 0 || !ok) { remove( tmp.c_str() ); return false; }
	return rename( tmp.c_str(), path.c_str() ) == 0; // readers see either the old file or the complete new one
}

} // namespace b200
