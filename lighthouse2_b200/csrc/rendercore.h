// rendercore.h -- RenderCore_B200: device state + host logic behind the CoreAPI.
// Mirrors the role of lib/RenderCore_Bart/rendercore.h:21-49 and lib/rendercore_optix7/rendercore.h.
#pragma once
#include "kernels.h"
#include "wide_bvh.h"
#include "../../include/rendercore_b200.h"
#include <string>
#include <vector>

namespace b200 {

// Growable device array (cudaMalloc); grow() preserves contents with a device-to-device copy.
struct DeviceBuffer {
	void* ptr = nullptr;
	size_t bytes = 0;
	cudaError_t reserve( size_t need, bool keep, cudaStream_t s );
	void release();
};

// growable page-locked host array (uploads and downloads that run inside a frame or beside kernels must not be pageable)
struct PinnedBuffer {
	void* ptr = nullptr;
	size_t bytes = 0;
	cudaError_t reserve( size_t need )
	{
		if (need <= bytes) return cudaSuccess;
		if (ptr) cudaFreeHost( ptr );
		ptr = nullptr, bytes = 0;
		const size_t cap = (need * 2 + 4095) & ~(size_t)4095;
		cudaError_t e = cudaHostAlloc( &ptr, cap, cudaHostAllocDefault );
		if (e == cudaSuccess) bytes = cap;
		return e;
	}
	void release() { if (ptr) cudaFreeHost( ptr ); ptr = nullptr, bytes = 0; }
};

struct MeshRec {
	uint32_t nodeBase = 0, nodeCount = 0, triBase = 0, triCount = 0;
	std::vector<uint32_t> levelStart; // relative to nodeBase
	Aabb bounds;                      // object-space bounds of the current vertex positions
	bool built = false;
	DeviceBuffer refitLeafInfo; // (primitive, shading word) per leaf-order triangle, see refit.cu::leafSlotBox
	DeviceBuffer refitStash; // slot boxes of the nodes with inner children (16 float4 each), see refit.cu::refitWalkKernel
	DeviceBuffer refitParents, refitArrived; // per node: parent | inner children of the parent << 28; arrival counter (refit.cu::refitWalkKernel)
};

struct InstanceHost { int mesh = -1; float transform[16]; float inv[16]; double sigmaMin = 1; /* smallest singular value of inv's 3x3 */ };

struct WavefrontState; // wavefront.cu

class RenderCore {
public:
	explicit RenderCore( int device );
	~RenderCore();
	bool ok() const { return initOk; }

	// CoreAPI surface (lib/RenderSystem/core_api_base.h:84-113)
	int SetGeometry( int meshIdx, const float* vertexData, int vertexCount, int triangleCount, const void* coreTris, const uint32_t* alphaFlags );
	int SetInstance( int instanceIdx, int modelIdx, const float* transform );
	int UpdateToplevel();
	int Setting( const char* name, float value );
	int SettingWavefront( const char* name, float value ); // wavefront.cu: maxPathLength, spp, tile sharding ...
	int SetTarget( int width, int height, int spp );
	int SetPointLight( const float* pos, const float* radiance );
	int SetSky( const float* radiance );
	int SetLights( const void* area, int nArea, const void* point, int nPoint, const void* spot, int nSpot, const void* dir, int nDir ); // reference layouts
	int SetSkyData( const float* rgb, int width, int height ); // equirectangular float3 image
	int SetMaterials( const void* coreMaterials, int count ); // CoreMaterial[count], 688 B each; keeps color.value and roughness.value
	int SetMaterialColors( const float* rgba, int count );    // headless shortcut: (albedo rgb, roughness) per material; rgb > 1 = emissive
	int SetBlueNoise( const uint32_t* words, int count );     // the reference's 5 x 65536-word blue-noise table, or null
	int DisneyProbe( int mode, int n, const float* in28, float* out8 ); // diagnostic: the Disney BSDF alone (wavefront.cu::disneyProbeKernel)
	struct MaterialMap { int32_t texture; float scale, uvscale[2], uvoffset[2]; }; // per material: detail colour, normals, detail normals, roughness
	int uploadMaterials( const float* props16, const uint32_t* disneyWords, const MaterialMap* maps, int count );
	int SetMaterialTable( const float* props16, int count );  // 16 floats per material, see wavefront.cu
	int SetTextures( const void* coreTexDescs, int count );   // CoreTexDesc[count], host layout, 40 B each; ARGB32 maps are sampled
	int SetAccumulatorBuffer( void* devicePtr );              // caller-owned float4[w*h] accumulator (multi-GPU gathers), null = own
	int PeerHandle( void* out128 );                           // multi-GPU gather over peer memory: see wavefront.cu
	int OpenPeers( const void* handles, int world, int rank );
	int ClosePeers();
	int GatherFinalize( uint32_t totalSamples );
	long long peerLimitTicks() const;
	// Setting l2Persist: fraction of the L2 set aside for persisting lines (0 = off); streams that run wave kernels then carry an access
	// policy window over the node pool (the top of the tree is what every ray touches).  For scenes that do not fit the 126 MB L2.
	int applyL2Window( cudaStream_t s );
	float l2Persist = 0.f;
	int Finalize( uint32_t totalSamples );                    // frame = accumulator / totalSamples (after a gather)
	int Render( const void* viewPyramid, int converge );
	int ReadAccumulator( float* dst );
	int ReadFrame( float* dst );
	int ReadFrameAsync( float* dst );
	int WaitFrame();
	int ReadFilterBuffer( int which, void* dst ); // the per-pixel buffers of the last filtered frame (wavefront.cu)
	void* DeviceAccumulator();
	int SetProbe( int x, int y );

	// raw waves
	int TraceClosest( const float* O4, const float* D4, int n, float* hits );
	int TraceAny( const float* O4, const float* D4, int n, uint32_t* bits );
	int TraceClosestAsync( const float* O4, const float* D4, int n, float* hits ); // queued: see rendercore.cpp
	int TraceAnyAsync( const float* O4, const float* D4, int n, uint32_t* bits );
	int WaitWaves();
	int TraceClosestNoAlloc( const float* O4, const float* D4, int n, float* hits ); // b200_trace_closest_noalloc: see traverse.cu
	int TraceClosestDevice( const void* dO4, const void* dD4, int n, void* dHits, const uint32_t* dCount = nullptr );
	int TraceAnyDevice( const void* dO4, const void* dD4, int n, void* dBits, const uint32_t* dCount = nullptr );
	int TraceConnectDevice( const void* dO4, const void* dD4, int n, const void* dE4, void* dAccumulator, const uint32_t* dCount = nullptr,
		cudaStream_t onStream = nullptr, int counterSlot = 0 );
	int ReadTree( int which, void* dst, size_t bytes );
	int ReadStepCounters( int n, uint32_t* nodes, uint32_t* tris );
	int Refit( int meshIdx, const float* vertexData, int vertexCount );
	int RefitDevice( int meshIdx, const void* dVerts, int vertexCount, bool boundsKnown );

	int SetStream( void* s ) { stream = (cudaStream_t)s; return 0; }
	void GetStats( b200_stats* out );
	const std::string& Error() const { return error; }

	// shared with wavefront.cu
	int fail( const char* what, cudaError_t e );
	int fail( const std::string& what );
	TraceArgs baseArgs() const;
	cudaStream_t stream = 0;
	int smCount = 148;
	IsectMode isect = ISECT_REFERENCE;
	b200_stats stats{};
	DeviceBuffer fetchCounter; // one uint32 per wave launch slot
	float epsilon = 1e-4f;

private:
	int ensureToplevel();
	int hostWave( const float* O4, const float* D4, int n, void* out, bool anyHit, bool async );
	int buildOnDevice( int meshIdx, const float* vertexData, uint32_t T, const void* coreTris, double t0 );
	int uploadVertexNormals( const MeshRec& m, const void* coreTris );
	static constexpr int MAX_CHUNKS = 12;
	cudaStream_t copyIn = nullptr, copyOut = nullptr, stream2 = nullptr;
	cudaEvent_t chunkEv[2 * MAX_CHUNKS + 1] = {};
	// staging slots of the host-buffer waves: a wave queued with the *_async calls takes the next slot, so that its upload can run
	// while the previous wave is still being traced and downloaded; a slot is reused once the wave that had it is done
#ifndef B200_WAVE_SLOTS
#define B200_WAVE_SLOTS 3
#endif
	static constexpr int WAVE_SLOTS = B200_WAVE_SLOTS;
	cudaEvent_t slotDone[WAVE_SLOTS] = {};
	bool slotUsed[WAVE_SLOTS] = {};
	uint32_t waveSeq = 0;
	bool initOk = false;
	std::string error;
	int device = 0;
	// acceleration structure pools
	DeviceBuffer nodePool, triPool, instPool, shadePool, vnormPool, instShadePool; // vnormPool: per-triangle shading attributes (TRI_ATTR_STRIDE float4: vertex normals, uv, LOD, tangent, vertex alphas), only when CoreTris were supplied
	uint32_t blasNodeEnd = 0, triEnd = 0, tlasRoot = 0, tlasNodeCount = 0;
	float tlasTScale = 1.0f; // TraceArgs::tlasTScale of the current top level (UpdateToplevel)
	std::vector<MeshRec> meshes;
	std::vector<InstanceHost> instances;
	bool tlasDirty = true;
	bool deviceTlas = true;   // Setting deviceTlas: the top level is built on the device (build.cu::tlasBuildKernel) when it fits one thread block
	DeviceBuffer tlasIn, tlasScratch, tlasInfo;
	PinnedBuffer tlasStage;   // TlasInstanceInput[N], then 4 info words
	int updateToplevelDevice( uint32_t N, double t0 );
	// scratch
	DeviceBuffer rayO[WAVE_SLOTS], rayD[WAVE_SLOTS], hitBuf[WAVE_SLOTS], bitBuf[WAVE_SLOTS], vertStage, nodeLo, nodeHi, stepNodes, stepTris;
	bool refitWalk = true; // Setting refitWalk: the refit is one launch that climbs the tree (1) or one launch per level (0)
	int prepareRefit( MeshRec& m );
	bool countSteps = false, allowRefit = true;
	void* buildScratch = nullptr; size_t buildScratchBytes = 0; // scratch of the device builder, kept between builds once a mesh has been REbuilt (dynamic scenes), freed after the build otherwise
	bool keepBuildScratch = false;
	int gpuBuild = 4; // new meshes get their tree built on the device (build.cu): 4 = binned SAH from the top (default), 3 = SAH inside the cells of a Morton radix tree, 1 = the radix tree itself (LBVH), 2 = agglomerative clustering (PLOC); 0 = the host builder (bvh_builder.cpp)
	bool zeroCopy = true, twoStreams = true; int pinnedChunks = 1, zeroCopyBelow = 400000; // pinned host ray planes are fetched by the kernel itself (Setting zeroCopy / pinnedChunks)
	bool singleInstanceShortcut = true; // a one-instance scene skips the top-level node (Setting singleInstanceShortcut)
	int hostChunks = 4, hostChunkMode = 0; // chunk schedule of the host-buffer waves (Setting hostChunks / hostChunkMode)
	int buildThreads = 0;
	cudaEvent_t evA = nullptr, evB = nullptr;
	WavefrontState* wf = nullptr;
	friend struct WavefrontState;
};

void DestroyWavefront( WavefrontState* wf ); // wavefront.cu

// mat4::Inverted (lib/RenderSystem/common_types.h:568-611): cofactor inverse times 1/det, identity if det == 0.
void Mat4Inverted( const float* cell, float* out );

} // namespace b200
