// rendercore.cpp -- host side of RenderCore_B200 (scene upload, pools, wave launches, stats).
// Counterpart of lib/RenderCore_Bart/rendercore.cpp:47-101 (SetGeometry/SetInstance) and of
// lib/rendercore_optix7/rendercore.cpp:338-428 (SetGeometry/SetInstance/UpdateToplevel) for the B200 core.
#include "rendercore.h"
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <thread>

namespace b200 {

static double nowMs() { return std::chrono::duration<double, std::milli>( std::chrono::steady_clock::now().time_since_epoch() ).count(); }

// ---- mat4::Inverted (lib/RenderSystem/common_types.h:568-611) -----------------------------------------
// Cofactor expansion; each cofactor is six signed triple products accumulated left to right, which is the
// evaluation order of the reference expression, so the result is bit-identical for any matrix.
void Mat4Inverted( const float* c, float* out )
{
	static const int8_t T[16][6][4] = {
		{ { 1, 5, 10, 15 }, { -1, 5, 11, 14 }, { -1, 9, 6, 15 }, { 1, 9, 7, 14 }, { 1, 13, 6, 11 }, { -1, 13, 7, 10 } },
		{ { -1, 1, 10, 15 }, { 1, 1, 11, 14 }, { 1, 9, 2, 15 }, { -1, 9, 3, 14 }, { -1, 13, 2, 11 }, { 1, 13, 3, 10 } },
		{ { 1, 1, 6, 15 }, { -1, 1, 7, 14 }, { -1, 5, 2, 15 }, { 1, 5, 3, 14 }, { 1, 13, 2, 7 }, { -1, 13, 3, 6 } },
		{ { -1, 1, 6, 11 }, { 1, 1, 7, 10 }, { 1, 5, 2, 11 }, { -1, 5, 3, 10 }, { -1, 9, 2, 7 }, { 1, 9, 3, 6 } },
		{ { -1, 4, 10, 15 }, { 1, 4, 11, 14 }, { 1, 8, 6, 15 }, { -1, 8, 7, 14 }, { -1, 12, 6, 11 }, { 1, 12, 7, 10 } },
		{ { 1, 0, 10, 15 }, { -1, 0, 11, 14 }, { -1, 8, 2, 15 }, { 1, 8, 3, 14 }, { 1, 12, 2, 11 }, { -1, 12, 3, 10 } },
		{ { -1, 0, 6, 15 }, { 1, 0, 7, 14 }, { 1, 4, 2, 15 }, { -1, 4, 3, 14 }, { -1, 12, 2, 7 }, { 1, 12, 3, 6 } },
		{ { 1, 0, 6, 11 }, { -1, 0, 7, 10 }, { -1, 4, 2, 11 }, { 1, 4, 3, 10 }, { 1, 8, 2, 7 }, { -1, 8, 3, 6 } },
		{ { 1, 4, 9, 15 }, { -1, 4, 11, 13 }, { -1, 8, 5, 15 }, { 1, 8, 7, 13 }, { 1, 12, 5, 11 }, { -1, 12, 7, 9 } },
		{ { -1, 0, 9, 15 }, { 1, 0, 11, 13 }, { 1, 8, 1, 15 }, { -1, 8, 3, 13 }, { -1, 12, 1, 11 }, { 1, 12, 3, 9 } },
		{ { 1, 0, 5, 15 }, { -1, 0, 7, 13 }, { -1, 4, 1, 15 }, { 1, 4, 3, 13 }, { 1, 12, 1, 7 }, { -1, 12, 3, 5 } },
		{ { -1, 0, 5, 11 }, { 1, 0, 7, 9 }, { 1, 4, 1, 11 }, { -1, 4, 3, 9 }, { -1, 8, 1, 7 }, { 1, 8, 3, 5 } },
		{ { -1, 4, 9, 14 }, { 1, 4, 10, 13 }, { 1, 8, 5, 14 }, { -1, 8, 6, 13 }, { -1, 12, 5, 10 }, { 1, 12, 6, 9 } },
		{ { 1, 0, 9, 14 }, { -1, 0, 10, 13 }, { -1, 8, 1, 14 }, { 1, 8, 2, 13 }, { 1, 12, 1, 10 }, { -1, 12, 2, 9 } },
		{ { -1, 0, 5, 14 }, { 1, 0, 6, 13 }, { 1, 4, 1, 14 }, { -1, 4, 2, 13 }, { -1, 12, 1, 6 }, { 1, 12, 2, 5 } },
		{ { 1, 0, 5, 10 }, { -1, 0, 6, 9 }, { -1, 4, 1, 10 }, { 1, 4, 2, 9 }, { 1, 8, 1, 6 }, { -1, 8, 2, 5 } } };
	float inv[16];
	for (int i = 0; i < 16; i++)
	{
		float acc = 0;
		for (int k = 0; k < 6; k++)
		{
			const int8_t* t = T[i][k];
			volatile float first = t[0] < 0 ? -c[t[1]] : c[t[1]];
			volatile float p = first * c[t[2]];
			volatile float q = p * c[t[3]];
			acc = k == 0 ? (float)q : acc + q;
		}
		inv[i] = acc;
	}
	const float det = c[0] * inv[0] + c[1] * inv[4] + c[2] * inv[8] + c[3] * inv[12];
	for (int i = 0; i < 16; i++) out[i] = (i % 5 == 0) ? 1.f : 0.f;
	if (det != 0)
	{
		const float invdet = 1.0f / det;
		for (int i = 0; i < 16; i++) out[i] = inv[i] * invdet;
	}
}

// smallest singular value of the upper 3x3 of a row-major 4x4: sqrt of the smallest eigenvalue of A^T A (symmetric 3x3, closed form)
static double SmallestSingularValue3x3( const float* v )
{
	double B[3][3];
	for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) B[r][c] = (double)v[r] * v[c] + (double)v[4 + r] * v[4 + c] + (double)v[8 + r] * v[8 + c];
	const double q = (B[0][0] + B[1][1] + B[2][2]) / 3.0;
	const double p1 = B[0][1] * B[0][1] + B[0][2] * B[0][2] + B[1][2] * B[1][2];
	const double p2 = (B[0][0] - q) * (B[0][0] - q) + (B[1][1] - q) * (B[1][1] - q) + (B[2][2] - q) * (B[2][2] - q) + 2.0 * p1;
	double lmin = q;
	if (p2 > 1e-300)
	{
		const double pp = std::sqrt( p2 / 6.0 );
		double C[3][3];
		for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) C[r][c] = (B[r][c] - (r == c ? q : 0.0)) / pp;
		double det = C[0][0] * (C[1][1] * C[2][2] - C[1][2] * C[2][1]) - C[0][1] * (C[1][0] * C[2][2] - C[1][2] * C[2][0]) + C[0][2] * (C[1][0] * C[2][1] - C[1][1] * C[2][0]);
		det = std::min( 1.0, std::max( -1.0, det * 0.5 ) );
		lmin = q + 2.0 * pp * std::cos( std::acos( det ) / 3.0 + 2.0943951023931953 ); // the smallest of the three eigenvalues
	}
	return std::sqrt( std::max( lmin, 0.0 ) );
}

// ---- DeviceBuffer -------------------------------------------------------------------------------------
cudaError_t DeviceBuffer::reserve( size_t need, bool keep, cudaStream_t s )
{
	if (need <= bytes) return cudaSuccess;
	size_t cap = bytes ? bytes : 256;
	while (cap < need) cap += cap / 2 + 256;
	cap = (cap + 255) & ~(size_t)255;
	void* np = nullptr;
	cudaError_t e = cudaMalloc( &np, cap );
	if (e != cudaSuccess) return e;
	if (ptr)
	{
		if (keep)
		{
			e = cudaMemcpyAsync( np, ptr, bytes, cudaMemcpyDeviceToDevice, s );
			if (e == cudaSuccess) e = cudaStreamSynchronize( s );
			if (e != cudaSuccess) { cudaFree( np ); return e; }
		}
		else cudaStreamSynchronize( s );
		cudaFree( ptr );
	}
	ptr = np, bytes = cap;
	return cudaSuccess;
}

void DeviceBuffer::release()
{
	if (ptr) cudaFree( ptr );
	ptr = nullptr, bytes = 0;
}

// ---- lifetime -----------------------------------------------------------------------------------------
RenderCore::RenderCore( int dev ) : device( dev )
{
	int count = 0;
	cudaError_t e = cudaGetDeviceCount( &count );
	if (e != cudaSuccess || count == 0) { fail( "no CUDA device: RenderCore_B200 has no CPU fallback", e ); return; }
	if (dev < 0 || dev >= count) { fail( "device index out of range" ); return; }
	if ((e = cudaSetDevice( dev )) != cudaSuccess) { fail( "cudaSetDevice", e ); return; }
	cudaDeviceProp prop;
	if ((e = cudaGetDeviceProperties( &prop, dev )) != cudaSuccess) { fail( "cudaGetDeviceProperties", e ); return; }
	if (prop.major < 10) { fail( std::string( "RenderCore_B200 kernels are built for sm_100a only; found " ) + prop.name ); return; }
	smCount = prop.multiProcessorCount;
	stats.smCount = smCount, stats.ccMajor = prop.major, stats.ccMinor = prop.minor;
	stats.vramMB = (uint32_t)(prop.totalGlobalMem >> 20);
	stats.probedInstid = stats.probedTriid = -1;
	if ((e = cudaEventCreate( &evA )) != cudaSuccess || (e = cudaEventCreate( &evB )) != cudaSuccess) { fail( "cudaEventCreate", e ); return; }
	if ((e = fetchCounter.reserve( 256, false, 0 )) != cudaSuccess) { fail( "cudaMalloc", e ); return; }
	buildThreads = (int)std::thread::hardware_concurrency();
	buildThreads = buildThreads < 1 ? 1 : (buildThreads > 64 ? 64 : buildThreads); // (the passes are memory-bound well before 64)
	initOk = true;
}

RenderCore::~RenderCore()
{
	if (!initOk) return;
	cudaSetDevice( device );
	cudaDeviceSynchronize();
	DestroyWavefront( wf );
	DeviceBuffer* all[] = { &fetchCounter, &nodePool, &triPool, &instPool, &shadePool, &vnormPool, &instShadePool, &vertStage, &nodeLo, &nodeHi,
		&stepNodes, &stepTris, &tlasIn, &tlasScratch, &tlasInfo };
	for (DeviceBuffer* b : all) b->release();
	for (MeshRec& m : meshes) m.refitParents.release(), m.refitArrived.release(), m.refitStash.release(), m.refitLeafInfo.release();
	tlasStage.release();
	if (buildScratch) cudaFree( buildScratch );
	for (int i = 0; i < WAVE_SLOTS; i++) rayO[i].release(), rayD[i].release(), hitBuf[i].release(), bitBuf[i].release();
	if (evA) cudaEventDestroy( evA );
	if (evB) cudaEventDestroy( evB );
	if (copyIn) { cudaStreamDestroy( copyIn ), cudaStreamDestroy( copyOut ), cudaStreamDestroy( stream2 ); for (cudaEvent_t ev : chunkEv) if (ev) cudaEventDestroy( ev ); for (cudaEvent_t ev : slotDone) if (ev) cudaEventDestroy( ev ); }
}

int RenderCore::fail( const char* what, cudaError_t e )
{
	error = std::string( what ) + ": " + cudaGetErrorString( e );
	return 1;
}
int RenderCore::fail( const std::string& what )
{
	error = what;
	return 1;
}

int RenderCore::Setting( const char* name, float value )
{
	// keys sent every frame by RenderSystem::Render (lib/RenderSystem/rendersystem.cpp:231-236)
	if (!strcmp( name, "epsilon" )) { epsilon = value; return 0; }
	// ("clampDirect", "clampIndirect", "filter", "TAA": with the frame settings in wavefront.cu -- honoured when Setting filterCore is on)
	// B200-specific keys
	if (!strcmp( name, "isectMode" )) { isect = value >= 0.5f ? ISECT_WATERTIGHT : ISECT_REFERENCE; return 0; }
	if (!strcmp( name, "gpuBuild" )) { gpuBuild = value >= 3.5f ? 4 : (value >= 2.5f ? 3 : (value >= 1.5f ? 2 : (value >= 0.5f ? 1 : 0))); return 0; } // 1: radix tree, 2: PLOC, 3: SAH inside the radix tree's cells, 4: binned SAH from the top (build.cu)
	if (!strcmp( name, "deviceTlas" )) { deviceTlas = value >= 0.5f; tlasDirty = true; return 0; }
	if (!strcmp( name, "twoStreams" )) { twoStreams = value >= 0.5f; return 0; }
	if (!strcmp( name, "zeroCopy" )) { zeroCopy = value >= 0.5f; return 0; }
	if (!strcmp( name, "pinnedChunks" )) { pinnedChunks = (int)value; return 0; }
	if (!strcmp( name, "zeroCopyBelow" )) { zeroCopyBelow = (int)value; return 0; }
	if (!strcmp( name, "hostChunks" )) { hostChunks = (int)value; return 0; }
	if (!strcmp( name, "hostChunkMode" )) { hostChunkMode = (int)value; return 0; }
	if (!strcmp( name, "singleInstanceShortcut" )) { singleInstanceShortcut = value >= 0.5f; return 0; }
	if (!strcmp( name, "countSteps" )) { countSteps = value >= 0.5f; return 0; }
	if (!strcmp( name, "refit" )) { allowRefit = value >= 0.5f; return 0; }
	if (!strcmp( name, "refitWalk" )) { refitWalk = value >= 0.5f; return 0; }
	if (!strcmp( name, "l2Persist" )) { if (value < 0 || value > 1) return fail( "Setting l2Persist: 0 (off) .. 1 (fraction of the L2 that may be set aside)" ); l2Persist = value; return 0; }
	if (!strcmp( name, "gpuBuildLevels" )) { if (!(value >= 1)) return fail( "Setting gpuBuildLevels: >= 1" ); g_sahLevels = value > 64 ? 64 : (int)value; return 0; }
	if (!strcmp( name, "sahPrimCost" )) { if (!(value > 0)) return fail( "Setting sahPrimCost: > 0" ); g_sahPrimCost = value; return 0; }
	if (!strcmp( name, "buildThreads" )) { buildThreads = value < 1 ? 1 : (int)value; return 0; }
	return SettingWavefront( name, value );
}


// ---- geometry -----------------------------------------------------------------------------------------
// shading record per ORIGINAL primitive: geometric normal (CoreTri::Nx,Ny,Nz when supplied, common_classes.h:69-71,
// else from the vertices) + CoreTri::material (:67); what the shade stage reads via instanceTriangles[PRIMIDX]
// in the reference (lib/rendercore_optix7/kernels/pathtracer.h:70-100)
struct ShadeRec { float n[3]; uint32_t material; };
static void makeShadeRecords( const float* vertexData, const uint32_t T, const void* coreTris, std::vector<ShadeRec>& shade, int threads )
{
	shade.resize( T );
	ParallelRange( threads, T, 65536, [&]( size_t begin, size_t end ) { for (size_t i = begin; i < end; i++)
	{
		const float *p0 = vertexData + (size_t)i * 12, *p1 = p0 + 4, *p2 = p0 + 8;
		const float e1[3] = { p1[0] - p0[0], p1[1] - p0[1], p1[2] - p0[2] }, e2[3] = { p2[0] - p0[0], p2[1] - p0[1], p2[2] - p0[2] };
		float n[3] = { e1[1] * e2[2] - e1[2] * e2[1], e1[2] * e2[0] - e1[0] * e2[2], e1[0] * e2[1] - e1[1] * e2[0] };
		const float l2 = n[0] * n[0] + n[1] * n[1] + n[2] * n[2], il = l2 > 0 ? 1.0f / sqrtf( l2 ) : 0.f;
		shade[i].n[0] = n[0] * il, shade[i].n[1] = n[1] * il, shade[i].n[2] = n[2] * il, shade[i].material = 0;
		if (coreTris)
		{
			const char* ct = (const char*)coreTris + (size_t)i * 176;
			float N[3];
			memcpy( &N[0], ct + 44, 4 ), memcpy( &N[1], ct + 60, 4 ), memcpy( &N[2], ct + 76, 4 );
			if (N[0] * N[0] + N[1] * N[1] + N[2] * N[2] > 0.5f) shade[i].n[0] = N[0], shade[i].n[1] = N[1], shade[i].n[2] = N[2];
			memcpy( &shade[i].material, ct + 28, 4 );
		}
	} } );
}

// per-triangle shading attributes of a mesh from its CoreTris (common_classes.h:57-91): vertex normals vN0/vN1/vN2 (offsets 32/48/64) for
// SMOOTH materials, texture coordinates u0..2 (0) / v0..2 (16) and LOD (124) for mapped ones.  4 float4 per ORIGINAL primitive, indexed
// like the shading records: (vN0,u0) (vN1,u1) (vN2,u2) (v0,v1,v2,LOD).  The pool exists only once a mesh has supplied CoreTris; other
// meshes read zeros (= flat, unmapped).
int RenderCore::uploadVertexNormals( const MeshRec& m, const void* coreTris )
{
	if (!coreTris && !vnormPool.ptr) return 0;
	cudaError_t e;
	constexpr size_t REC = TRI_ATTR_STRIDE * 16; // bytes per triangle
	const size_t oldBytes = vnormPool.bytes, need = (size_t)(m.triBase + m.triCount + 1) * REC;
	if ((e = vnormPool.reserve( need, true, stream )) != cudaSuccess) return fail( "triangle attribute pool", e );
	if (vnormPool.bytes > oldBytes && (e = cudaMemsetAsync( (char*)vnormPool.ptr + oldBytes, 0, vnormPool.bytes - oldBytes, stream )) != cudaSuccess) return fail( "triangle attribute pool", e );
	char* dst = (char*)vnormPool.ptr + (size_t)m.triBase * REC;
	if (!coreTris) e = cudaMemsetAsync( dst, 0, (size_t)m.triCount * REC, stream );
	else
	{
		// CoreTri (common_classes.h:62-83): u0..2 @0, v0..2 @16, vN0 @32, vN1 @48, vN2 @64, T @80, B @96, alpha @112 (w = LOD)
		std::vector<float> at( (size_t)m.triCount * TRI_ATTR_STRIDE * 4, 0.f );
		ParallelRange( buildThreads, m.triCount, 65536, [&]( size_t begin, size_t end ) { for (size_t i = begin; i < end; i++)
		{
			const char* ct = (const char*)coreTris + (size_t)i * 176;
			float* o = &at[(size_t)i * TRI_ATTR_STRIDE * 4];
			memcpy( o, ct + 32, 12 ), memcpy( o + 4, ct + 48, 12 ), memcpy( o + 8, ct + 64, 12 );
			memcpy( o + 3, ct, 4 ), memcpy( o + 7, ct + 4, 4 ), memcpy( o + 11, ct + 8, 4 );
			memcpy( o + 12, ct + 16, 12 ), memcpy( o + 15, ct + 124, 4 );
			memcpy( o + 16, ct + 80, 12 ), memcpy( o + 19, ct + 112, 4 ), memcpy( o + 20, ct + 96, 12 ), memcpy( o + 23, ct + 116, 4 ), memcpy( o + 24, ct + 120, 4 );
			// CoreTri::ltriIdx (@12; -1 = not a light, pathtracer.h:139) kept + 1, so that the zero-filled records of meshes without CoreTris mean "unknown"
			int32_t ltri;
			memcpy( &ltri, ct + 12, 4 );
			ltri = ltri >= 0 ? ltri + 1 : 0;
			memcpy( o + 25, &ltri, 4 );
		} } );
		e = cudaMemcpyAsync( dst, at.data(), at.size() * 4, cudaMemcpyHostToDevice, stream );
		if (e == cudaSuccess) e = cudaStreamSynchronize( stream ); // the staging vector dies here
	}
	return e == cudaSuccess ? 0 : fail( "triangle attribute upload", e );
}

// SetGeometry with Setting gpuBuild >= 1 (the default): the vertices go up once, the topology is built on the device (build.cu) and the refit
// that follows fills boxes, quantised planes and triangle records.  Nothing but 12 bytes per level comes back to the host.
int RenderCore::buildOnDevice( int meshIdx, const float* vertexData, uint32_t T, const void* coreTris, double t0 )
{
	MeshRec m;
	m.nodeBase = blasNodeEnd, m.triBase = triEnd, m.triCount = T;
	if ((uint64_t)m.triBase + T + 32 >= (1ull << 27)) return fail( "SetGeometry: more than 2^27 triangles in one core (queue entries carry 27-bit triangle indices)" );
	// wide nodes: a seventh to a tenth of the triangles in practice, one per binary inner node at worst: room for T / 2 first, for all of them if the collapse runs out
	uint32_t maxNodes = T / 2 + 64;
	// shading records: the refit below writes the geometric normals; only CoreTri::material has to come from the host
	std::vector<ShadeRec> shade;
	if (coreTris)
	{
		shade.resize( T );
		for (uint32_t i = 0; i < T; i++)
		{
			shade[i].n[0] = shade[i].n[1] = shade[i].n[2] = 0;
			memcpy( &shade[i].material, (const char*)coreTris + (size_t)i * 176 + 28, 4 );
		}
	}
	cudaError_t e;
	if ((e = nodePool.reserve( (size_t)(m.nodeBase + maxNodes + 64) * 80, true, stream )) != cudaSuccess) return fail( "node pool", e );
	if ((e = triPool.reserve( (size_t)(m.triBase + T + 1) * 48, true, stream )) != cudaSuccess) return fail( "triangle pool", e );
	if ((e = shadePool.reserve( (size_t)(m.triBase + T + 1) * 16, true, stream )) != cudaSuccess) return fail( "shading pool", e );
	if ((e = vertStage.reserve( (size_t)T * 48 + 16, false, stream )) != cudaSuccess) return fail( "vertex staging", e );
	if ((e = cudaMemcpyAsync( vertStage.ptr, vertexData, (size_t)T * 48, cudaMemcpyHostToDevice, stream )) != cudaSuccess) return fail( "vertex upload", e );
	if (coreTris) e = cudaMemcpyAsync( (char*)shadePool.ptr + (size_t)m.triBase * 16, shade.data(), (size_t)T * 16, cudaMemcpyHostToDevice, stream );
	else e = cudaMemsetAsync( (char*)shadePool.ptr + (size_t)m.triBase * 16, 0, (size_t)T * 16, stream );
	if (e != cudaSuccess) return fail( "shading upload", e );
	// B200_BUILD_TIMING=1: where the time of a device build goes (each mark waits for the stream)
	static const bool timing = getenv( "B200_BUILD_TIMING" ) != nullptr;
	double tMark = nowMs();
	auto mark = [&]( const char* what ) { if (!timing) return; cudaStreamSynchronize( stream ); const double t = nowMs(); fprintf( stderr, "device build, %u triangles: %s %.2f ms\n", T, what, t - tMark ); tMark = t; };
	if (timing) fprintf( stderr, "device build, %u triangles: host side before the upload %.2f ms\n", T, tMark - t0 );
	mark( "pools + upload" );
	e = GpuBuildTopology( (const float4*)vertStage.ptr, T, (uint4*)((char*)nodePool.ptr + (size_t)m.nodeBase * 80), (float4*)((char*)triPool.ptr + (size_t)m.triBase * 48),
		m.nodeBase, m.triBase, maxNodes, stream, m.levelStart, m.nodeCount, gpuBuild, &buildScratch, &buildScratchBytes );
	mark( "topology" );
	if (e == cudaErrorMemoryAllocation)
	{
		cudaGetLastError();
		maxNodes = T + 64;
		if ((e = nodePool.reserve( (size_t)(m.nodeBase + maxNodes + 64) * 80, true, stream )) != cudaSuccess) return fail( "node pool", e );
		e = GpuBuildTopology( (const float4*)vertStage.ptr, T, (uint4*)((char*)nodePool.ptr + (size_t)m.nodeBase * 80), (float4*)((char*)triPool.ptr + (size_t)m.triBase * 48),
			m.nodeBase, m.triBase, maxNodes, stream, m.levelStart, m.nodeCount, gpuBuild, &buildScratch, &buildScratchBytes );
	}
	if (e != cudaSuccess) return fail( "GPU BVH build", e );
	if ((int)m.levelStart.size() - 1 > MAX_TREE_LEVELS) return fail( "SetGeometry: the device-built tree is deeper than the traversal stack allows (wide_bvh.h MAX_TREE_LEVELS)" );
	if (prepareRefit( m )) return 1;
	mark( "refit links" );
	m.built = true;
	blasNodeEnd = m.nodeBase + m.nodeCount, triEnd = m.triBase + T;
	meshes.push_back( m );
	if (uploadVertexNormals( m, coreTris )) return 1;
	if (RefitDevice( meshIdx, vertStage.ptr, (int)(3 * T), false )) return 1; // boxes, planes, triangle records; fetches the root box
	mark( "attributes + refit" );
	if (!keepBuildScratch && buildScratch) cudaStreamSynchronize( stream ), cudaFree( buildScratch ), buildScratch = nullptr, buildScratchBytes = 0;
	stats.bvhBuildTime += (float)(nowMs() - t0);
	stats.wideNodeCount = blasNodeEnd, stats.triangleCount = triEnd;
	return 0;
}


// CoreAPI_Base::SetGeometry (core_api_base.h:108).  vertexData is float4[3T], triangle-expanded.  The first
// call for a mesh builds its wide BVH on the host and appends it to the pools; later calls with the same
// triangle count refit on the GPU (the reference instead rebuilds: rendercore.cpp:78).
int RenderCore::SetGeometry( int meshIdx, const float* vertexData, int vertexCount, int triangleCount, const void* coreTris, const uint32_t* )
{
	if (meshIdx < 0 || triangleCount < 0 || vertexCount != 3 * triangleCount) return fail( "SetGeometry: bad arguments (vertexCount must be 3*triangleCount)" );
	if (triangleCount > 0 && !vertexData) return fail( "SetGeometry: null vertexData" );
	if (meshIdx > (int)meshes.size()) return fail( "SetGeometry: mesh indices must arrive densely" );
	if (meshIdx < (int)meshes.size() && meshes[meshIdx].built)
	{
		if ((uint32_t)triangleCount == meshes[meshIdx].triCount && allowRefit)
		{
			if (coreTris && uploadVertexNormals( meshes[meshIdx], coreTris )) return 1; // animated meshes bring new vertex normals
			return Refit( meshIdx, vertexData, vertexCount );
		}
		if ((uint32_t)triangleCount != meshes[meshIdx].triCount && meshIdx != (int)meshes.size() - 1)
			return fail( "SetGeometry: changing the triangle count is only supported for the most recently added mesh" );
		if (meshIdx != (int)meshes.size() - 1 && !allowRefit)
			return fail( "SetGeometry: rebuild (refit=0) is only supported for the most recently added mesh" );
		// rebuild of the last mesh: drop it and fall through to a fresh build (a scene that does this once will do it again: the builder's scratch stays)
		keepBuildScratch = true;
		blasNodeEnd = meshes[meshIdx].nodeBase, triEnd = meshes[meshIdx].triBase;
		meshes.back().refitParents.release(), meshes.back().refitArrived.release(), meshes.back().refitStash.release(), meshes.back().refitLeafInfo.release();
		meshes.pop_back();
	}
	const double t0 = nowMs();
	const uint32_t T = (uint32_t)triangleCount;
	// on-disk cache of the host-built tree (bvh_cache.cpp), keyed by the vertex data; B200_BVH_CACHE names the directory
	static const char* cacheEnv = getenv( "B200_BVH_CACHE" );
	const bool useCache = cacheEnv && *cacheEnv && T >= 1024;
	// the tree is built on the device (Setting gpuBuild, default 4: binned SAH from the top) unless the mesh is tiny, the host builder
	// was asked for (gpuBuild = 0) or its trees are kept on disk
	if (gpuBuild && T >= 64 && !useCache) return buildOnDevice( meshIdx, vertexData, T, coreTris, t0 );
	// (the passes over all triangles of a large mesh -- here, in the builder and in the staging below -- run on buildThreads threads)
	std::vector<Aabb> boxes( T );
	ParallelRange( buildThreads, T, 65536, [&]( size_t begin, size_t end ) { for (size_t i = begin; i < end; i++)
	{
		boxes[i].reset();
		for (int k = 0; k < 3; k++) boxes[i].grow( vertexData + (size_t)(3 * i + k) * 4 );
	} } );
	WideBvh bvh;
	uint64_t cacheKey = 0;
	bool fromCache = false;
	if (useCache)
	{
		cacheKey = HashBytes( vertexData, (size_t)T * 48, 0xB200 );
		fromCache = LoadWideBvh( cacheEnv, cacheKey, T, 3, bvh );
	}
	if (!fromCache)
	{
		BuildWideBvh( boxes.data(), T, 3, buildThreads, bvh );
		if (useCache) StoreWideBvh( cacheEnv, cacheKey, T, 3, bvh ); // best effort: an unwritable directory only costs the next start-up
	}
	else stats.bvhCacheHits++;
	if ((int)bvh.levelStart.size() - 1 > MAX_TREE_LEVELS) return fail( "SetGeometry: the tree is deeper than the traversal stack allows (wide_bvh.h MAX_TREE_LEVELS)" );
	MeshRec m;
	m.nodeBase = blasNodeEnd, m.nodeCount = (uint32_t)bvh.nodes.size(), m.triBase = triEnd, m.triCount = T;
	m.levelStart = bvh.levelStart, m.bounds = bvh.bounds, m.built = true;
	for (WideNode& n : bvh.nodes) n.childBase += m.nodeBase, n.primBase += m.triBase; // absolute pool indices
	if ((uint64_t)m.triBase + T + 32 >= (1ull << 27)) return fail( "SetGeometry: more than 2^27 triangles in one core (queue entries carry 27-bit triangle indices)" );
	std::vector<WideTri> tris( T );
	ParallelRange( buildThreads, T, 65536, [&]( size_t begin, size_t end ) { for (size_t j = begin; j < end; j++)
	{
		const uint32_t prim = bvh.primOrder[j];
		const float *p0 = vertexData + (size_t)prim * 12, *p1 = p0 + 4, *p2 = p0 + 8;
		WideTri& t = tris[j];
		t.prim = prim, t.pad1 = t.pad2 = 0;
		for (int a = 0; a < 3; a++) t.v0[a] = p0[a], t.e1[a] = p1[a] - p0[a], t.e2[a] = p2[a] - p0[a];
	} } );
	std::vector<ShadeRec> shade;
	makeShadeRecords( vertexData, T, coreTris, shade, buildThreads );
	cudaError_t e;
	if ((e = nodePool.reserve( (size_t)(m.nodeBase + m.nodeCount + 64) * 80, true, stream )) != cudaSuccess) return fail( "node pool", e );
	if ((e = triPool.reserve( (size_t)(m.triBase + T + 1) * 48, true, stream )) != cudaSuccess) return fail( "triangle pool", e );
	if ((e = shadePool.reserve( (size_t)(m.triBase + T + 1) * 16, true, stream )) != cudaSuccess) return fail( "shading pool", e );
	if ((e = cudaMemcpyAsync( (char*)nodePool.ptr + (size_t)m.nodeBase * 80, bvh.nodes.data(), (size_t)m.nodeCount * 80, cudaMemcpyHostToDevice, stream )) != cudaSuccess) return fail( "node upload", e );
	if (T && (e = cudaMemcpyAsync( (char*)triPool.ptr + (size_t)m.triBase * 48, tris.data(), (size_t)T * 48, cudaMemcpyHostToDevice, stream )) != cudaSuccess) return fail( "triangle upload", e );
	if (T && (e = cudaMemcpyAsync( (char*)shadePool.ptr + (size_t)m.triBase * 16, shade.data(), (size_t)T * 16, cudaMemcpyHostToDevice, stream )) != cudaSuccess) return fail( "shading upload", e );
	if (prepareRefit( m )) return 1;
	if ((e = cudaStreamSynchronize( stream )) != cudaSuccess) return fail( "SetGeometry sync", e ); // the staging vectors die here
	blasNodeEnd = m.nodeBase + m.nodeCount, triEnd = m.triBase + T;
	meshes.push_back( m );
	if (uploadVertexNormals( m, coreTris )) return 1;
	tlasDirty = true;
	stats.bvhBuildTime += (float)(nowMs() - t0);
	stats.wideNodeCount = blasNodeEnd, stats.triangleCount = triEnd;
	return 0;
}

// b200_refit: new positions for an existing mesh (host float4[3T])
int RenderCore::Refit( int meshIdx, const float* vertexData, int vertexCount )
{
	if (meshIdx < 0 || meshIdx >= (int)meshes.size() || !meshes[meshIdx].built) return fail( "Refit: unknown mesh" );
	MeshRec& m = meshes[meshIdx];
	if ((uint32_t)vertexCount != 3 * m.triCount) return fail( "Refit: vertex count does not match the built mesh" );
	cudaError_t e;
	if ((e = vertStage.reserve( (size_t)vertexCount * 16 + 16, false, stream )) != cudaSuccess) return fail( "vertex staging", e );
	if ((e = cudaMemcpyAsync( vertStage.ptr, vertexData, (size_t)vertexCount * 16, cudaMemcpyHostToDevice, stream )) != cudaSuccess) return fail( "vertex upload", e );
	// object-space bounds for the top level come from the host copy of the vertices (no read-back needed)
	Aabb b;
	b.reset();
	for (int i = 0; i < vertexCount; i++) b.grow( vertexData + (size_t)i * 4 );
	m.bounds = b;
	return RefitDevice( meshIdx, vertStage.ptr, vertexCount, true );
}

// parent links and arrival counters of a freshly placed tree, for the single-launch refit
int RenderCore::prepareRefit( MeshRec& m )
{
	cudaError_t e;
	// refitParents = parent links | stash numbers | their count (one buffer); the stash holds the eight slot boxes of every node with inner children
	const size_t n1 = (size_t)m.nodeCount + 1;
	if ((e = m.refitParents.reserve( (2 * n1 + 1) * 4, false, stream )) != cudaSuccess || (e = m.refitArrived.reserve( n1 * 4, false, stream )) != cudaSuccess)
		return fail( "refit links", e );
	uint32_t* links = (uint32_t*)m.refitParents.ptr;
	if ((e = LaunchRefitParents( (const uint4*)nodePool.ptr, m.nodeBase, m.nodeCount, links, (uint32_t*)m.refitArrived.ptr, links + n1, links + 2 * n1, stream )) != cudaSuccess)
		return fail( "refit links", e );
	stats.kernelLaunches++;
	uint32_t innerNodes = 0;
	if ((e = cudaMemcpyAsync( &innerNodes, links + 2 * n1, 4, cudaMemcpyDeviceToHost, stream )) != cudaSuccess || (e = cudaStreamSynchronize( stream )) != cudaSuccess)
		return fail( "refit links", e );
	if ((e = m.refitStash.reserve( ((size_t)innerNodes + 1) * 256, false, stream )) != cudaSuccess) return fail( "refit stash", e );
	// leaf-order copy of what the refit needs of the records it rewrites (the triangle records carry the primitive indices, the shading records the material words, by now)
	if ((e = m.refitLeafInfo.reserve( ((size_t)m.triCount + 1) * 8, false, stream )) != cudaSuccess) return fail( "refit leaf info", e );
	if ((e = LaunchRefitLeafInfo( (const float4*)triPool.ptr, (const float4*)shadePool.ptr, m.triBase, m.triCount, (uint2*)m.refitLeafInfo.ptr, stream )) != cudaSuccess) return fail( "refit leaf info", e );
	stats.kernelLaunches++;
	return 0;
}

int RenderCore::RefitDevice( int meshIdx, const void* dVerts, int vertexCount, bool boundsKnown )
{
	if (meshIdx < 0 || meshIdx >= (int)meshes.size() || !meshes[meshIdx].built) return fail( "Refit: unknown mesh" );
	MeshRec& m = meshes[meshIdx];
	if ((uint32_t)vertexCount != 3 * m.triCount) return fail( "Refit: vertex count does not match the built mesh" );
	cudaError_t e;
	if ((e = nodeLo.reserve( (size_t)m.nodeCount * 16, false, stream )) != cudaSuccess) return fail( "refit scratch", e );
	if ((e = nodeHi.reserve( (size_t)m.nodeCount * 16, false, stream )) != cudaSuccess) return fail( "refit scratch", e );
	RefitArgs a;
	a.nodes = (uint4*)nodePool.ptr, a.tris = (float4*)triPool.ptr, a.verts = (const float4*)dVerts;
	a.nodeLo = (float4*)nodeLo.ptr, a.nodeHi = (float4*)nodeHi.ptr, a.shade = (float4*)shadePool.ptr;
	a.nodeBase = m.nodeBase, a.triBase = m.triBase, a.triCount = m.triCount;
	a.leafInfo = (const uint2*)m.refitLeafInfo.ptr;
	cudaEventRecord( evA, stream );
	if (refitWalk && m.refitParents.ptr)
	{
		// one launch: every group does its node's triangles, the last to arrive at a node finishes it and climbs
		const uint32_t* links = (const uint32_t*)m.refitParents.ptr;
		if ((e = LaunchRefitWalk( a, m.nodeCount, links, (uint32_t*)m.refitArrived.ptr, links + m.nodeCount + 1, (float4*)m.refitStash.ptr, stream )) != cudaSuccess) return fail( "refit", e );
		stats.kernelLaunches++;
	}
	else
	{
	// deepest level first; the tiny top levels (at most 16 nodes each) share one single-CTA launch
	const int levels = (int)m.levelStart.size() - 1;
	int nTop = 0;
	while (nTop < levels && nTop < 16 && m.levelStart[nTop + 1] - m.levelStart[nTop] <= 16) nTop++; // measured: a single CTA is slower than one launch per level beyond a few nodes
	for (int level = levels - 1; level >= nTop; level--)
	{
		if ((e = LaunchRefitLevel( a, m.levelStart[level], m.levelStart[level + 1], stream )) != cudaSuccess) return fail( "refit level", e );
		stats.kernelLaunches++;
	}
	if (nTop)
	{
		if ((e = LaunchRefitTop( a, m.levelStart.data(), nTop, stream )) != cudaSuccess) return fail( "refit top levels", e );
		stats.kernelLaunches++;
	}
	}
	cudaEventRecord( evB, stream );
	if (!boundsKnown)
	{
		// caller supplied device positions only: fetch the refitted root box (32 bytes)
		float lo[4], hi[4];
		if ((e = cudaMemcpyAsync( lo, nodeLo.ptr, 16, cudaMemcpyDeviceToHost, stream )) != cudaSuccess) return fail( "root box", e );
		if ((e = cudaMemcpyAsync( hi, nodeHi.ptr, 16, cudaMemcpyDeviceToHost, stream )) != cudaSuccess) return fail( "root box", e );
		if ((e = cudaStreamSynchronize( stream )) != cudaSuccess) return fail( "refit sync", e );
		for (int c = 0; c < 3; c++) m.bounds.lo[c] = lo[c], m.bounds.hi[c] = hi[c];
	}
	else if ((e = cudaStreamSynchronize( stream )) != cudaSuccess) return fail( "refit sync", e );
	float ms = 0;
	cudaEventElapsedTime( &ms, evA, evB );
	stats.refitTime = ms;
	tlasDirty = true;
	return 0;
}

// ---- instances ----------------------------------------------------------------------------------------
// CoreAPI_Base::SetInstance (core_api_base.h:110); modelIdx == -1 ends the stream and truncates
// (lib/RenderSystem/rendersystem.cpp:167-169, lib/RenderCore_Bart/rendercore.cpp:84-89).
int RenderCore::SetInstance( int instanceIdx, int modelIdx, const float* transform )
{
	if (instanceIdx < 0) return fail( "SetInstance: negative index" );
	if (modelIdx == -1)
	{
		if ((int)instances.size() > instanceIdx) instances.resize( instanceIdx ), tlasDirty = true;
		return 0;
	}
	if (modelIdx < 0 || modelIdx >= (int)meshes.size()) return fail( "SetInstance: unknown mesh" );
	if (instanceIdx > (int)instances.size()) return fail( "SetInstance: instance indices must arrive densely" );
	if (instanceIdx == (int)instances.size()) instances.emplace_back();
	InstanceHost& in = instances[instanceIdx];
	in.mesh = modelIdx;
	static const float ident[16] = { 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1 };
	memcpy( in.transform, transform ? transform : ident, 64 );
	Mat4Inverted( in.transform, in.inv );
	in.sigmaMin = SmallestSingularValue3x3( in.inv );
	tlasDirty = true;
	return 0;
}

// CoreAPI_Base::UpdateToplevel (core_api_base.h:112): full top-level rebuild, every frame if anything moved
// (the CPU core rebuilds per Render, rendercore.cpp:199; optix7 rebuilds the IAS, rendercore.cpp:387-428).
int RenderCore::UpdateToplevel()
{
	const double t0 = nowMs();
	const uint32_t N = (uint32_t)instances.size();
	// Hit distances are kept in the object-space units of the instance that was hit (the ray direction is re-normalised at every
	// instance, bvh.cpp:607-610, and the reference compares those numbers across instances).  A hit at world distance d inside instance k
	// has the object-space distance d * |inv_k D| >= d * sigma_min( inv_k ), so a top-level box entered at world distance d can be culled
	// against the current hit distance t only when d * min_k sigma_min( inv_k ) > t.  For rigid instances the factor is 1.
	{
		double smin = 1e300;
		for (uint32_t i = 0; i < N; i++) smin = std::min( smin, instances[i].sigmaMin );
		// (a margin for the float arithmetic of the instance entry; a degenerate instance switches top-level culling off)
		tlasTScale = (N && smin > 1e-12) ? (float)std::min( 1.0 / (smin * (1.0 - 1.0 / 4096.0)), 1e30 ) : (N ? 1e30f : 1.0f);
	}
	if (deviceTlas && N >= 1 && N <= MAX_DEVICE_TLAS_INSTANCES) return updateToplevelDevice( N, t0 );
	std::vector<Aabb> boxes( N );
	for (uint32_t i = 0; i < N; i++)
	{
		// world box = all 8 transformed corners of the mesh box (the CPU reference transforms only two,
		// bvh.cpp:579-587, which is exact for translations and too small under rotation)
		const Aabb& mb = meshes[instances[i].mesh].bounds;
		const float* c = instances[i].transform;
		boxes[i].reset();
		if (meshes[instances[i].mesh].triCount == 0) { float z[3] = { c[3], c[7], c[11] }; boxes[i].grow( z ); continue; }
		for (int k = 0; k < 8; k++)
		{
			const float x = (k & 1) ? mb.hi[0] : mb.lo[0], y = (k & 2) ? mb.hi[1] : mb.lo[1], z = (k & 4) ? mb.hi[2] : mb.lo[2];
			float p[3] = { c[0] * x + c[1] * y + c[2] * z + c[3], c[4] * x + c[5] * y + c[6] * z + c[7], c[8] * x + c[9] * y + c[10] * z + c[11] };
			boxes[i].grow( p );
		}
		// one ulp-scale pad: the device transforms the RAY instead of the box, so roundings differ
		for (int a = 0; a < 3; a++)
		{
			const float pad = 4e-7f * fmaxf( fabsf( boxes[i].lo[a] ), fabsf( boxes[i].hi[a] ) ) + 1e-30f;
			boxes[i].lo[a] -= pad, boxes[i].hi[a] += pad;
		}
	}
	WideBvh tlas;
	BuildWideBvh( boxes.data(), N, 1, 1, tlas );
	{
		// the kernels' per-lane stack has no bound check: refuse scenes that could overrun it (see wide_bvh.h)
		int blasLevels = 0;
		for (uint32_t i = 0; i < N; i++) blasLevels = std::max( blasLevels, (int)meshes[instances[i].mesh].levelStart.size() - 1 );
		if (blasLevels + 2 * ((int)tlas.levelStart.size() - 1) > TRAVERSAL_STACK_ENTRIES)
			return fail( "UpdateToplevel: bottom-level + 2 x top-level tree depth exceeds the traversal stack (wide_bvh.h TRAVERSAL_STACK_ENTRIES)" );
	}
	std::vector<InstanceRec> recs( N );
	for (uint32_t j = 0; j < N; j++)
	{
		const uint32_t id = tlas.primOrder[j];
		InstanceRec& r = recs[j];
		memset( &r, 0, sizeof( r ) );
		memcpy( r.inv, instances[id].inv, 64 );
		r.blasRoot = meshes[instances[id].mesh].nodeBase, r.instId = id, r.meshId = (uint32_t)instances[id].mesh;
		const float* v = r.inv;
		r.flags = (v[12] == 0 && v[13] == 0 && v[14] == 0 && v[15] == 1) ? 1u : 0u;
	}
	// shading-time instance table by INSTANCE ID (the reference's CoreInstanceDesc array, rendercore_optix7/rendercore.cpp:639-674):
	// forward transform rows (normals to world space, pathtracer.h via GetShadingData) + the mesh's first triangle in the pools
	struct InstShade { float row[3][4]; uint32_t triBase, meshId, pad0, pad1; };
	std::vector<InstShade> ish( N );
	for (uint32_t i = 0; i < N; i++)
	{
		memcpy( ish[i].row, instances[i].transform, 48 );
		ish[i].triBase = meshes[instances[i].mesh].triBase, ish[i].meshId = (uint32_t)instances[i].mesh, ish[i].pad0 = ish[i].pad1 = 0;
	}
	tlasRoot = blasNodeEnd, tlasNodeCount = (uint32_t)tlas.nodes.size();
	for (WideNode& n : tlas.nodes) n.childBase += tlasRoot; // primBase indexes the instance records directly
	cudaError_t e;
	if ((e = nodePool.reserve( (size_t)(tlasRoot + tlasNodeCount) * 80, true, stream )) != cudaSuccess) return fail( "node pool", e );
	if ((e = instPool.reserve( (size_t)(N + 1) * 96, false, stream )) != cudaSuccess) return fail( "instance pool", e );
	if ((e = cudaMemcpyAsync( (char*)nodePool.ptr + (size_t)tlasRoot * 80, tlas.nodes.data(), (size_t)tlasNodeCount * 80, cudaMemcpyHostToDevice, stream )) != cudaSuccess) return fail( "tlas upload", e );
	if (N && (e = cudaMemcpyAsync( instPool.ptr, recs.data(), (size_t)N * 96, cudaMemcpyHostToDevice, stream )) != cudaSuccess) return fail( "instance upload", e );
	if ((e = instShadePool.reserve( (size_t)(N + 1) * 64, false, stream )) != cudaSuccess) return fail( "instance shading pool", e );
	if (N && (e = cudaMemcpyAsync( instShadePool.ptr, ish.data(), (size_t)N * 64, cudaMemcpyHostToDevice, stream )) != cudaSuccess) return fail( "instance shading upload", e );
	if ((e = cudaStreamSynchronize( stream )) != cudaSuccess) return fail( "UpdateToplevel sync", e );
	tlasDirty = false;
	stats.tlasTime = (float)(nowMs() - t0);
	stats.instanceCount = N, stats.tlasNodeCount = tlasNodeCount;
	return 0;
}

// The top level built by one thread block on the device (build.cu::tlasBuildKernel): the host hands over the instances' matrices and
// mesh bounds (176 bytes each, from a pinned staging array) and reads back four words (node count, level count, overflow flag).
int RenderCore::updateToplevelDevice( uint32_t N, double t0 )
{
	cudaError_t e;
	const uint32_t maxNodes = N + 8;
	if ((e = tlasStage.reserve( (size_t)N * sizeof( TlasInstanceInput ) + 64 )) != cudaSuccess) return fail( "top-level staging", e );
	TlasInstanceInput* in = (TlasInstanceInput*)tlasStage.ptr;
	uint32_t* hostInfo = (uint32_t*)((char*)tlasStage.ptr + (size_t)N * sizeof( TlasInstanceInput ));
	int blasLevels = 0;
	for (uint32_t i = 0; i < N; i++)
	{
		const InstanceHost& h = instances[i];
		const MeshRec& m = meshes[h.mesh];
		TlasInstanceInput& t = in[i];
		memcpy( t.transform, h.transform, 64 ), memcpy( t.inv, h.inv, 64 );
		for (int a = 0; a < 3; a++) t.lo[a] = m.bounds.lo[a], t.hi[a] = m.bounds.hi[a];
		t.blasRoot = m.nodeBase, t.triBase = m.triBase, t.meshId = (uint32_t)h.mesh;
		t.flags = (h.inv[12] == 0 && h.inv[13] == 0 && h.inv[14] == 0 && h.inv[15] == 1) ? 1u : 0u;
		t.empty = m.triCount == 0 ? 1u : 0u, t.pad = 0;
		blasLevels = std::max( blasLevels, (int)m.levelStart.size() - 1 );
	}
	tlasRoot = blasNodeEnd;
	if ((e = nodePool.reserve( (size_t)(tlasRoot + maxNodes + 64) * 80, true, stream )) != cudaSuccess) return fail( "node pool", e );
	if ((e = instPool.reserve( (size_t)(N + 1) * 96, false, stream )) != cudaSuccess) return fail( "instance pool", e );
	if ((e = instShadePool.reserve( (size_t)(N + 1) * 64, false, stream )) != cudaSuccess) return fail( "instance shading pool", e );
	if ((e = tlasIn.reserve( (size_t)N * sizeof( TlasInstanceInput ), false, stream )) != cudaSuccess || (e = tlasScratch.reserve( TlasScratchBytes( N ), false, stream )) != cudaSuccess ||
		(e = tlasInfo.reserve( 64, false, stream )) != cudaSuccess) return fail( "top-level buffers", e );
	if ((e = cudaMemcpyAsync( tlasIn.ptr, in, (size_t)N * sizeof( TlasInstanceInput ), cudaMemcpyHostToDevice, stream )) != cudaSuccess) return fail( "instance upload", e );
	if ((e = LaunchTlasBuild( (const TlasInstanceInput*)tlasIn.ptr, N, (uint4*)((char*)nodePool.ptr + (size_t)tlasRoot * 80), tlasRoot, maxNodes, (float4*)instPool.ptr,
		(float4*)instShadePool.ptr, (uint32_t*)tlasInfo.ptr, tlasScratch.ptr, 1.0f, stream )) != cudaSuccess) return fail( "top-level build kernel", e );
	stats.kernelLaunches++;
	if ((e = cudaMemcpyAsync( hostInfo, tlasInfo.ptr, 64, cudaMemcpyDeviceToHost, stream )) != cudaSuccess) return fail( "top-level info", e );
	if ((e = cudaStreamSynchronize( stream )) != cudaSuccess) return fail( "UpdateToplevel sync", e );
	static const bool tlasTiming = getenv( "B200_TLAS_TIMING" ) != nullptr; // debugging aid: the kernel's phase clock (SM cycles) on stderr
	if (tlasTiming)
		fprintf( stderr, "[tlas] n=%u nodes=%u levels=%u cycles: boxes %u keys %u sort %u tree %u bottom-up %u collapse %u planes %u records %u\n", N, hostInfo[0], hostInfo[1],
			hostInfo[4], hostInfo[5], hostInfo[6], hostInfo[7], hostInfo[8], hostInfo[9], hostInfo[10], hostInfo[11] );
	if (hostInfo[2]) return fail( "UpdateToplevel: the device top-level build overflowed its node budget" );
	tlasNodeCount = hostInfo[0];
	// the kernels' per-lane stack has no bound check: refuse scenes that could overrun it (see wide_bvh.h)
	if (blasLevels + 2 * (int)hostInfo[1] > TRAVERSAL_STACK_ENTRIES)
		return fail( "UpdateToplevel: bottom-level + 2 x top-level tree depth exceeds the traversal stack (wide_bvh.h TRAVERSAL_STACK_ENTRIES)" );
	tlasDirty = false;
	stats.tlasTime = (float)(nowMs() - t0);
	stats.instanceCount = N, stats.tlasNodeCount = tlasNodeCount;
	return 0;
}

int RenderCore::ensureToplevel() { return tlasDirty ? UpdateToplevel() : 0; }


// Setting l2Persist: an access-policy window over the node pool on stream `s` (see rendercore.h).  The set-aside is a device-wide limit.
int RenderCore::applyL2Window( cudaStream_t s )
{
	cudaStreamAttrValue attr;
	memset( &attr, 0, sizeof( attr ) );
	if (l2Persist <= 0 || !nodePool.ptr)
	{
		attr.accessPolicyWindow.num_bytes = 0; // no window
		cudaStreamSetAttribute( s, cudaStreamAttributeAccessPolicyWindow, &attr );
		cudaGetLastError();
		return 0;
	}
	int maxPersist = 0, maxWindow = 0;
	cudaDeviceGetAttribute( &maxPersist, cudaDevAttrMaxPersistingL2CacheSize, device );
	cudaDeviceGetAttribute( &maxWindow, cudaDevAttrMaxAccessPolicyWindowSize, device );
	const size_t setAside = (size_t)((double)maxPersist * l2Persist);
	cudaError_t e = cudaDeviceSetLimit( cudaLimitPersistingL2CacheSize, setAside );
	if (e != cudaSuccess) return fail( "cudaLimitPersistingL2CacheSize", e );
	const size_t poolBytes = (size_t)(blasNodeEnd + tlasNodeCount) * 80;
	const size_t window = std::min( poolBytes, (size_t)maxWindow );
	attr.accessPolicyWindow.base_ptr = nodePool.ptr;
	attr.accessPolicyWindow.num_bytes = window;
	attr.accessPolicyWindow.hitRatio = window ? (float)std::min( 1.0, (double)setAside / (double)window ) : 0.f;
	attr.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
	attr.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
	if ((e = cudaStreamSetAttribute( s, cudaStreamAttributeAccessPolicyWindow, &attr )) != cudaSuccess) return fail( "access policy window", e );
	return 0;
}

// ---- the raw waves -------------------------------------------------------------------------------------
TraceArgs RenderCore::baseArgs() const
{
	TraceArgs a;
	memset( &a, 0, sizeof( a ) );
	a.nodes = (const uint4*)nodePool.ptr, a.tris = (const float4*)triPool.ptr, a.instances = (const float4*)instPool.ptr;
	a.fetchCounter = (uint32_t*)fetchCounter.ptr;
	a.tlasRoot = tlasRoot;
	a.singleInstance = (instances.size() == 1 && singleInstanceShortcut) ? 1u : 0u;
	a.magic = 0x4B000000u;
	a.tlasTScale = tlasTScale;
	return a;
}

int RenderCore::TraceClosestDevice( const void* dO4, const void* dD4, int n, void* dHits, const uint32_t* dCount )
{
	if (n < 0) return fail( "trace: negative ray count" );
	if (ensureToplevel()) return 1;
	if (n == 0) return 0;
	TraceArgs a = baseArgs();
	a.rayO = (const float4*)dO4, a.rayD = (const float4*)dD4, a.hits = (float4*)dHits, a.count = (uint32_t)n, a.countPtr = dCount;
	cudaError_t e;
	if (countSteps)
	{
		if ((e = stepNodes.reserve( (size_t)n * 4, false, stream )) != cudaSuccess) return fail( "step counters", e );
		if ((e = stepTris.reserve( (size_t)n * 4, false, stream )) != cudaSuccess) return fail( "step counters", e );
		a.stepNodes = (uint32_t*)stepNodes.ptr, a.stepTris = (uint32_t*)stepTris.ptr;
	}
	cudaEventRecord( evA, stream );
	if ((e = LaunchExtend( a, isect, countSteps, smCount, stream )) != cudaSuccess) return fail( "extend kernel launch", e );
	cudaEventRecord( evB, stream );
	stats.kernelLaunches++;
	return 0;
}

int RenderCore::TraceAnyDevice( const void* dO4, const void* dD4, int n, void* dBits, const uint32_t* dCount )
{
	if (n < 0) return fail( "trace: negative ray count" );
	if (ensureToplevel()) return 1;
	if (n == 0) return 0;
	TraceArgs a = baseArgs();
	a.rayO = (const float4*)dO4, a.rayD = (const float4*)dD4, a.occluded = (uint32_t*)dBits, a.count = (uint32_t)n, a.countPtr = dCount;
	cudaError_t e;
	if ((e = cudaMemsetAsync( dBits, 0, (size_t)((n + 31) / 32) * 4, stream )) != cudaSuccess) return fail( "clear bitmask", e );
	cudaEventRecord( evA, stream );
	if ((e = LaunchShadow( a, isect, smCount, stream )) != cudaSuccess) return fail( "shadow kernel launch", e );
	cudaEventRecord( evB, stream );
	stats.kernelLaunches++;
	return 0;
}

// shadow wave fused with the connect step: unoccluded rays add E4.xyz to accumulator[pixelIdx = bits of E4.w]
// (rendercore_optix7/optix/.optix.cu:137-149; RenderCore_OptixPrime_B/kernels/connections.h:23-34)
int RenderCore::TraceConnectDevice( const void* dO4, const void* dD4, int n, const void* dE4, void* dAccumulator, const uint32_t* dCount,
	cudaStream_t onStream, int counterSlot )
{
	cudaStream_t s = onStream ? onStream : stream;
	if (n < 0) return fail( "trace: negative ray count" );
	if (!dE4 || !dAccumulator) return fail( "connect: null potentials / accumulator" );
	if (ensureToplevel()) return 1;
	if (n == 0) return 0;
	TraceArgs a = baseArgs();
	a.rayO = (const float4*)dO4, a.rayD = (const float4*)dD4, a.potentials = (const float4*)dE4, a.accumulator = (float4*)dAccumulator;
	a.count = (uint32_t)n, a.countPtr = dCount;
	a.fetchCounter += counterSlot; // a wave that runs beside another one needs its own fetch counter
	cudaError_t e;
	if (!onStream) cudaEventRecord( evA, s );
	if ((e = LaunchShadow( a, isect, smCount, s )) != cudaSuccess) return fail( "connect kernel launch", e );
	if (!onStream) cudaEventRecord( evB, s );
	stats.kernelLaunches++;
	return 0;
}

// The host-buffer waves are pipelined: the rays go up in chunks on a copy-in stream, each chunk's kernel waits only for its
// own upload, and its results go down on a copy-out stream while the next chunk is traced (PCIe is full duplex, so upload,
// kernel and download overlap; with pinned host memory the call costs about the upload alone).
// Chunk schedule of a host-buffer wave.  mode 0: `parts` equal chunks (at least 256K rays each); mode 1: halving chunks
// (n/2, n/4, ... down to 32K rays) so that little work is left after the last byte has arrived.
static int chunkSchedule( int n, int* offsets, int maxChunks, int mode, int parts )
{
	int count = 0, off = 0;
	if (parts > maxChunks) parts = maxChunks;
	if (parts < 1) parts = 1;
	const int equal = (((n / parts > (1 << 18) ? n / parts : (1 << 18))) + 1023) & ~1023;
	while (off < n)
	{
		int m = mode ? ((n - off) / 2 + 1023) & ~1023 : equal;
		if (mode && m < (1 << 15)) m = n - off;
		if (count == maxChunks - 1 || (n - off) - m < (1 << 14)) m = n - off; // no tiny tail chunk
		offsets[count++] = off;
		off += m;
	}
	offsets[count] = n;
	return count;
}

int RenderCore::hostWave( const float* O4, const float* D4, int n, void* out, bool anyHit, bool async )
{
	if (n < 0) return fail( "trace: negative ray count" );
	if (n == 0) return ensureToplevel();
	if (ensureToplevel()) return 1;
	cudaError_t e;
	const size_t rb = (size_t)n * 16, bb = (size_t)((n + 31) / 32) * 4;
	// staging slot of this wave (see rendercore.h); growing a buffer frees the old one, which waits for the device
	const int slot = (int)(waveSeq++ % WAVE_SLOTS);
	DeviceBuffer &rayO = this->rayO[slot], &rayD = this->rayD[slot], &hitBuf = this->hitBuf[slot], &bitBuf = this->bitBuf[slot];
	DeviceBuffer& outBuf = anyHit ? bitBuf : hitBuf;
	if ((e = rayO.reserve( rb, false, stream )) != cudaSuccess || (e = rayD.reserve( rb, false, stream )) != cudaSuccess ||
		(e = outBuf.reserve( anyHit ? bb : rb, false, stream )) != cudaSuccess) return fail( "ray buffers", e );
	if (countSteps && ((e = stepNodes.reserve( (size_t)n * 4, false, stream )) != cudaSuccess || (e = stepTris.reserve( (size_t)n * 4, false, stream )) != cudaSuccess))
		return fail( "step counters", e );
	if (!copyIn)
	{
		if ((e = cudaStreamCreateWithFlags( &copyIn, cudaStreamNonBlocking )) != cudaSuccess || (e = cudaStreamCreateWithFlags( &copyOut, cudaStreamNonBlocking )) != cudaSuccess ||
			(e = cudaStreamCreateWithFlags( &stream2, cudaStreamNonBlocking )) != cudaSuccess) return fail( "copy streams", e );
		for (int i = 0; i < 2 * MAX_CHUNKS + 1; i++) if ((e = cudaEventCreateWithFlags( &chunkEv[i], cudaEventDisableTiming )) != cudaSuccess) return fail( "chunk events", e );
		for (int i = 0; i < WAVE_SLOTS; i++) if ((e = cudaEventCreateWithFlags( &slotDone[i], cudaEventDisableTiming )) != cudaSuccess) return fail( "slot events", e );
	}
	// Small waves over pinned (page-locked) ray planes: the kernel itself fetches the rays over PCIe, no staging copy and no
	// upload -> trace dependency.  (Measured: wins below ~0.5M rays; above, chunked DMA uploads overlap better.)
	const float4 *zcO = nullptr, *zcD = nullptr;
	if (zeroCopy && !async && !countSteps && n <= zeroCopyBelow) // (queued waves: the DMA uploads overlap the previous wave anyway; measured 4.03 vs 4.15 ms per step)
	{
		cudaPointerAttributes pa, pb;
		if (cudaPointerGetAttributes( &pa, O4 ) == cudaSuccess && cudaPointerGetAttributes( &pb, D4 ) == cudaSuccess &&
			pa.type == cudaMemoryTypeHost && pb.type == cudaMemoryTypeHost && pa.devicePointer && pb.devicePointer)
			zcO = (const float4*)pa.devicePointer, zcD = (const float4*)pb.devicePointer;
		cudaGetLastError();
	}
	// The wave is cut into chunks: upload k+1 (copy engine), trace k, download k-1 (second copy engine) run concurrently.  Chunk
	// kernels alternate between two streams, so that the tail of one (few long rays left) overlaps the start of the next.
	int offsets[MAX_CHUNKS + 1];
	int nChunks = 1;
	offsets[0] = 0, offsets[1] = n;
	if (!countSteps) nChunks = chunkSchedule( n, offsets, MAX_CHUNKS, hostChunkMode, zcO ? pinnedChunks : hostChunks );
	const int join = 2 * MAX_CHUNKS;
	static const bool traceChunks = getenv( "B200_TRACE_CHUNKS" ) != nullptr; // debugging aid: per-chunk timeline on stderr
	static cudaEvent_t tev[4 * MAX_CHUNKS + 1];
	if (traceChunks && !tev[0]) for (cudaEvent_t& t : tev) cudaEventCreate( &t );
	// the uploads may start once the wave that had this staging slot is done (its kernels may still read the planes); the first
	// wave after other work on the core's stream also waits for that
	cudaEventRecord( chunkEv[join], stream );
	if (slotUsed[slot]) cudaStreamWaitEvent( copyIn, slotDone[slot], 0 ), cudaStreamWaitEvent( stream, slotDone[slot], 0 );
	else cudaStreamWaitEvent( copyIn, chunkEv[join], 0 );
	cudaStreamWaitEvent( stream2, chunkEv[join], 0 );
	if (traceChunks) cudaEventRecord( tev[4 * MAX_CHUNKS], stream );
	for (int k = 0; k < nChunks; k++)
	{
		const int off = offsets[k], m = offsets[k + 1] - off;
		const size_t ob = (size_t)off * 16, mb = (size_t)m * 16;
		cudaStream_t s = (k & 1) && twoStreams ? stream2 : stream;
		TraceArgs a = baseArgs();
		a.fetchCounter += k; // concurrent launches: one fetch counter each
		a.rayO = (const float4*)rayO.ptr + off, a.rayD = (const float4*)rayD.ptr + off, a.count = (uint32_t)m;
		if (zcO) a.hostO = zcO + off, a.hostD = zcD + off, a.stageO = (float4*)rayO.ptr + off, a.stageD = (float4*)rayD.ptr + off;
		else
		{
			if ((e = cudaMemcpyAsync( (char*)rayO.ptr + ob, (const char*)O4 + ob, mb, cudaMemcpyHostToDevice, copyIn )) != cudaSuccess) return fail( "ray upload", e );
			if ((e = cudaMemcpyAsync( (char*)rayD.ptr + ob, (const char*)D4 + ob, mb, cudaMemcpyHostToDevice, copyIn )) != cudaSuccess) return fail( "ray upload", e );
			cudaEventRecord( chunkEv[MAX_CHUNKS + k], copyIn );
			cudaStreamWaitEvent( s, chunkEv[MAX_CHUNKS + k], 0 );
			if (traceChunks) cudaEventRecord( tev[4 * k], copyIn );
		}
		if (traceChunks) cudaEventRecord( tev[4 * k + 1], s );
		if (k == nChunks - 1) cudaEventRecord( evA, s );
		const void* dOut;
		size_t outOff, outBytes;
		if (anyHit)
		{
			outOff = (size_t)(off / 32) * 4, outBytes = (size_t)((m + 31) / 32) * 4; // chunks start at multiples of 1024 rays
			a.occluded = (uint32_t*)((char*)bitBuf.ptr + outOff), dOut = a.occluded;
			if ((e = cudaMemsetAsync( a.occluded, 0, outBytes, s )) != cudaSuccess) return fail( "clear bitmask", e );
			if ((e = LaunchShadow( a, isect, smCount, s )) != cudaSuccess) return fail( "shadow kernel launch", e );
		}
		else
		{
			outOff = ob, outBytes = mb;
			a.hits = (float4*)hitBuf.ptr + off, dOut = a.hits;
			if (countSteps) a.stepNodes = (uint32_t*)stepNodes.ptr, a.stepTris = (uint32_t*)stepTris.ptr;
			if ((e = LaunchExtend( a, isect, countSteps, smCount, s )) != cudaSuccess) return fail( "extend kernel launch", e );
		}
		if (k == nChunks - 1) cudaEventRecord( evB, s );
		stats.kernelLaunches++;
		cudaEventRecord( chunkEv[k], s );
		cudaStreamWaitEvent( copyOut, chunkEv[k], 0 );
		if (traceChunks) cudaEventRecord( tev[4 * k + 2], s );
		if ((e = cudaMemcpyAsync( (char*)out + outOff, dOut, outBytes, cudaMemcpyDeviceToHost, copyOut )) != cudaSuccess) return fail( "result download", e );
		if (traceChunks) cudaEventRecord( tev[4 * k + 3], copyOut );
	}
	if (nChunks > 1 && twoStreams) { cudaEventRecord( chunkEv[join], stream2 ); cudaStreamWaitEvent( stream, chunkEv[join], 0 ); }
	cudaEventRecord( slotDone[slot], copyOut ); // the downloads wait for their kernels: everything of this wave is done by then
	slotUsed[slot] = true;
	if (async) return 0; // b200_wait collects the results
	if ((e = cudaStreamSynchronize( copyOut )) != cudaSuccess) return fail( anyHit ? "shadow wave" : "extend wave", e );
	if ((e = cudaStreamSynchronize( stream )) != cudaSuccess) return fail( anyHit ? "shadow wave" : "extend wave", e );
	cudaEventElapsedTime( &stats.lastTraceMs, evA, evB ); // device time of the LAST chunk's kernel
	if (traceChunks)
	{
		fprintf( stderr, "[chunks%s] n=%d %s:", zcO ? " zc" : "", n, anyHit ? "any" : "closest" );
		for (int k = 0; k < nChunks; k++)
		{
			float t[4] = {};
			for (int j = zcO ? 1 : 0; j < 4; j++) cudaEventElapsedTime( &t[j], tev[4 * MAX_CHUNKS], tev[4 * k + j] );
			fprintf( stderr, " [%dK up %.0f k %.0f-%.0f dn %.0f]", (offsets[k + 1] - offsets[k]) >> 10, t[0] * 1e3f, t[1] * 1e3f, t[2] * 1e3f, t[3] * 1e3f );
		}
		fprintf( stderr, "\n" );
	}
	return 0;
}

// host-buffer variants: H2D, wave, D2H, synchronise (the e2e path of bench.py)
int RenderCore::TraceClosest( const float* O4, const float* D4, int n, float* hits ) { return hostWave( O4, D4, n, hits, false, false ); }
int RenderCore::TraceAny( const float* O4, const float* D4, int n, uint32_t* bits ) { return hostWave( O4, D4, n, bits, true, false ); }

// Queued variants: the call returns once the wave's copies and kernels are enqueued; WaitWaves() returns when every queued wave
// has delivered its results.  Waves queued back to back overlap: the upload of wave k+1 runs while wave k is traced and its
// hits go down (the shape of an OptiX Prime query with RTP_QUERY_HINT_ASYNC + rtpQueryFinish; the reference's Prime core
// issues its queries synchronously, lib/RenderCore_OptixPrime_B/rendercore.cpp:528,562,587).  The host buffers must stay
// valid and untouched until WaitWaves(); pageable buffers make the copies synchronous, so pin them (b200_host_alloc).
int RenderCore::TraceClosestAsync( const float* O4, const float* D4, int n, float* hits ) { return hostWave( O4, D4, n, hits, false, true ); }
int RenderCore::TraceAnyAsync( const float* O4, const float* D4, int n, uint32_t* bits ) { return hostWave( O4, D4, n, bits, true, true ); }
int RenderCore::WaitWaves()
{
	if (!copyOut) return 0;
	cudaError_t e;
	if ((e = cudaStreamSynchronize( copyOut )) != cudaSuccess || (e = cudaStreamSynchronize( stream2 )) != cudaSuccess || (e = cudaStreamSynchronize( stream )) != cudaSuccess)
		return fail( "queued waves", e );
	return 0;
}

// b200_trace_closest_noalloc: the extend wave over host buffers without any allocation.  No counterpart in the reference:
// nothing on this path calls cudaMalloc, cudaFree or a routine that allocates -- which is also why a
// stale top level is an error here instead of being rebuilt (UpdateToplevel grows pools).
int RenderCore::TraceClosestNoAlloc( const float* O4, const float* D4, int n, float* hits )
{
	if (n < 0) return fail( "b200_trace_closest_noalloc: negative ray count" );
	if ((uint32_t)n > NOALLOC_MAX_RAYS)
		return fail( "b200_trace_closest_noalloc: " + std::to_string( n ) + " rays exceed the fixed scratch of B200_NOALLOC_MAX_RAYS = " + std::to_string( NOALLOC_MAX_RAYS ) +
			" rays (split the wave, or use b200_trace_closest)" );
	if (n > 0 && (!O4 || !D4 || !hits)) return fail( "b200_trace_closest_noalloc: null buffer" );
	if (tlasDirty) return fail( "b200_trace_closest_noalloc: the top level is out of date; call b200_update_toplevel first (this entry point must not allocate)" );
	if (n == 0) return 0;
	const cudaError_t e = b200::TraceClosestNoAlloc( baseArgs(), isect, smCount, device, O4, D4, (uint32_t)n, hits, stream );
	if (e != cudaSuccess) return fail( "b200_trace_closest_noalloc", e );
	stats.kernelLaunches++;
	return 0;
}

int RenderCore::ReadStepCounters( int n, uint32_t* nodes, uint32_t* tris )
{
	if (!stepNodes.ptr || stepNodes.bytes < (size_t)n * 4) return fail( "ReadStepCounters: no counters (Setting countSteps=1, then trace)" );
	cudaError_t e;
	if ((e = cudaMemcpy( nodes, stepNodes.ptr, (size_t)n * 4, cudaMemcpyDeviceToHost )) != cudaSuccess) return fail( "counters", e );
	if ((e = cudaMemcpy( tris, stepTris.ptr, (size_t)n * 4, cudaMemcpyDeviceToHost )) != cudaSuccess) return fail( "counters", e );
	return 0;
}

int RenderCore::ReadTree( int which, void* dst, size_t bytes )
{
	// the mesh trees as they lie in HBM: 80-byte wide nodes [0, wideNodeCount) or 48-byte triangle records [0, triangleCount)
	const DeviceBuffer& pool = which == 0 ? nodePool : triPool;
	const size_t used = which == 0 ? (size_t)blasNodeEnd * 80 : (size_t)triEnd * 48;
	if (which < 0 || which > 1 || bytes > used || !pool.ptr) return fail( "ReadTree: which = 0 (nodes) | 1 (triangles), bytes within the used part of the pool" );
	const cudaError_t e = cudaMemcpy( dst, pool.ptr, bytes, cudaMemcpyDeviceToHost );
	return e == cudaSuccess ? 0 : fail( "ReadTree", e );
}

void RenderCore::GetStats( b200_stats* out )
{
	// lastTraceMs of the device-pointer waves is resolved lazily here (the events may still be in flight)
	if (evA && evB && cudaEventQuery( evB ) == cudaSuccess)
	{
		float ms = 0;
		if (cudaEventElapsedTime( &ms, evA, evB ) == cudaSuccess) stats.lastTraceMs = ms;
	}
	cudaGetLastError();
	*out = stats;
}

} // namespace b200
