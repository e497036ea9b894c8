/* include/rendercore_b200.h -- C ABI of libRenderCore_B200.so.
 *
 * Two groups of entry points:
 *  (1) the Lighthouse 2 plugin symbols the RenderSystem binds with dlsym
 *      (lib/RenderSystem/core_api_base.cpp:119-132; declared lib/RenderCore_Bart/core_api.h:63-64):
 *          CoreAPI_Base* CreateCore();   void DestroyCore();
 *      CreateCore returns the C++ object whose vtable is restated in include/lh2_abi.h.
 *  (2) a flat, headless `b200_*` layer (plain pointers and sizes, no C++ and no torch types) that
 *      exposes the same operations one-to-one plus the raw extend/shadow queries, so tests, the bench
 *      and foreign hosts can drive the core without a C++ vtable.  Each function cites the reference
 *      interface it stands for.
 *
 * All functions return 0 on success, non-zero on failure (message via b200_last_error); the C++
 * CoreAPI shim turns non-zero into the reference's fatal-error convention (platform/system.cpp:221).
 * There is NO CPU fallback: b200_create fails if no sm_100 device is usable.
 *
 * Ray / hit wire formats:
 *   ray   : float4 O4 = (Ox,Oy,Oz,tmin), float4 D4 = (Dx,Dy,Dz,tmax)   -- 32 B/ray, two SoA planes
 *           (OptiX Prime RAY_ORIGIN_TMIN_DIRECTION_TMAX, RenderCore_OptixPrime_B/rendercore.cpp:199;
 *            the optix7 core stores the same two float4 planes, rendercore_optix7/optix/.optix.cu:125-135)
 *   hit   : float4 {as_float(u16 | v16<<16), as_float(instIdx), as_float(primIdx), t} -- 16 B/ray,
 *           miss = prim -1, inst -1(0xffffffff), t = 1e34  (rendercore_optix7/optix/.optix.cu:119-122,183-186)
 *   shadow: 1 bit per ray, bit set = occluded (Prime HIT_BITMASK, RenderCore_OptixPrime_B/rendercore.cpp:203)
 */
#ifndef RENDERCORE_B200_H
#define RENDERCORE_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define B200_API __attribute__((visibility("default")))

typedef struct b200_core b200_core; /* opaque */

/* isect modes, Setting("isectMode", v) */
#define B200_ISECT_REFERENCE 0  /* bit-exact restatement of Ray::IntersectTriangle + TLAS-leaf ray transform (parity runs) */
#define B200_ISECT_WATERTIGHT 1 /* fused, scale-invariant Moeller-Trumbore with shared-edge consistent rounding */

/* ---- lifetime: CreateCore/Init/Shutdown (lib/RenderCore_Bart/core_api.cpp:20-47,64-69) ---- */
B200_API int b200_create( b200_core** out, int device );
B200_API void b200_destroy( b200_core* c );
B200_API const char* b200_last_error( void );
/* all work of this core is enqueued on `stream` (a cudaStream_t; 0 = the legacy default stream) */
B200_API int b200_set_stream( b200_core* c, void* stream );

/* ---- scene: CoreAPI_Base::SetGeometry / SetInstance / UpdateToplevel (core_api_base.h:108-112) ---- */
/* vertexData: float4[3*triangleCount], triangle-expanded (host_mesh.cpp:525-527).  triangles: optional
 * CoreTri[triangleCount] (176 B each; only `material` is kept).  First call for a meshIdx builds the
 * wide BVH; later calls with the same triangle count refit it on the GPU (Setting "refit" = 0 forces a rebuild). */
B200_API int b200_set_geometry( b200_core* c, int meshIdx, const float* vertexData, int vertexCount, int triangleCount,
	const void* triangles, const uint32_t* alphaFlags );
/* transform: row-major float[16] (mat4::cell).  modelIdx == -1 ends the stream and truncates
 * (lib/RenderSystem/rendersystem.cpp:167-169). */
B200_API int b200_set_instance( b200_core* c, int instanceIdx, int modelIdx, const float* transform );
B200_API int b200_update_toplevel( b200_core* c );
/* GPU refit of an already-built mesh from new positions (host float4[3T]); what SetGeometry does on update */
B200_API int b200_refit( b200_core* c, int meshIdx, const float* vertexData, int vertexCount );
/* same, positions already resident in device memory (device float4[3T]) */
B200_API int b200_refit_device( b200_core* c, int meshIdx, const void* dVertexData, int vertexCount );

/* ---- CoreAPI_Base::Setting (core_api_base.h:92; keys rendersystem.cpp:231-236 + isectMode, refit, countSteps, buildThreads,
 * maxPathLength, shardIndex, shardCount, sampleOffset, jitter, pathsPerPass) ---- */
B200_API int b200_setting( b200_core* c, const char* name, float value );

/* ---- raw wavefront queries: the extend / shadow waves ---- */
/* host buffers; H2D + kernel + D2H + synchronise inside the call */
B200_API int b200_trace_closest( b200_core* c, const float* O4, const float* D4, int n, float* hits );
B200_API int b200_trace_any( b200_core* c, const float* O4, const float* D4, int n, uint32_t* occludedBits );
/* queued variants: the call returns once the wave's copies and kernels are enqueued; b200_wait returns when every queued wave has
 * delivered its results.  Waves queued back to back overlap (upload of wave k+1 while wave k is traced and downloaded; up to
 * three in flight, then a wave waits for the staging slot of the wave three before it).  Host buffers: pinned (b200_host_alloc),
 * valid and untouched until b200_wait.  The shape of an OptiX Prime query with RTP_QUERY_HINT_ASYNC + rtpQueryFinish (the
 * reference issues its Prime queries synchronously, lib/RenderCore_OptixPrime_B/rendercore.cpp:528,562,587). */
B200_API int b200_trace_closest_async( b200_core* c, const float* O4, const float* D4, int n, float* hits );
B200_API int b200_trace_any_async( b200_core* c, const float* O4, const float* D4, int n, uint32_t* occludedBits );
B200_API int b200_wait( b200_core* c );
/* The extend wave over host buffers WITHOUT ANY ALLOCATION.  This entry point has no counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same inputs, outputs and results as b200_trace_closest (bit for bit), but the wave kernel's scratch -- its fetch counter
 * and the device staging of the two ray planes and the hit records -- is a file-scope static device array of fixed size (traverse.cu)
 * instead of the core's growable buffers, and nothing on the path calls cudaMalloc, cudaFree or an allocating routine.
 *   limit : n <= B200_NOALLOC_MAX_RAYS; a larger n is rejected with an error (split the wave or use b200_trace_closest).  The array
 *           (256 B + 48 B per ray = 3 MB) is held by every GPU that loads the library until the process exits.
 *   state : the scene must be built and its top level current (b200_update_toplevel) -- a stale top level is an error here, because
 *           rebuilding it may grow pools.
 *   races : all calls share the one array, which keeps the last call's contents (the call clears what must start at zero); two calls that
 *           overlap on the GPU -- other streams, other cores on the same device, other threads -- race.  The call synchronises its stream
 *           before it returns. */
#define B200_NOALLOC_MAX_RAYS 65536
B200_API int b200_trace_closest_noalloc( b200_core* c, const float* O4, const float* D4, int n, float* hits );
/* device buffers (cudaMalloc'ed by anyone in this process); asynchronous on the core's stream.
 * occludedBits must hold (n+31)/32 words; it is cleared by the call. */
B200_API int b200_trace_closest_device( b200_core* c, const void* dO4, const void* dD4, int n, void* dHits );
B200_API int b200_trace_any_device( b200_core* c, const void* dO4, const void* dD4, int n, void* dOccludedBits );
/* shadow wave fused with the connect step (.optix.cu:137-149, connections.h:23-34): every ray that is NOT occluded adds
 * E4[i].xyz to dAccumulator[pixel], pixel = integer bits of E4[i].w (atomicAdd; device buffers, asynchronous) */
B200_API int b200_trace_connect_device( b200_core* c, const void* dO4, const void* dD4, int n, const void* dE4, void* dAccumulator );
/* per-ray traversal counters of the LAST b200_trace_closest call with Setting("countSteps",1):
 * nodes visited and triangles tested (the analogue of the reference's `int* c`, bvh.cpp:259) */
B200_API int b200_read_step_counters( b200_core* c, int n, uint32_t* nodesVisited, uint32_t* trisTested );
/* the mesh trees as they lie in HBM (diagnostics; the reference keeps its BVH on the host, bvh.h:60-76): which = 0 the 80-byte wide
 * nodes [0, b200_stats.wideNodeCount), 1 the 48-byte triangle records [0, b200_stats.triangleCount); bytes <= the used part */
B200_API int b200_read_tree( b200_core* c, int which, void* dst, size_t bytes );

/* ---- frame: SetTarget / SetLights / Render / GetCoreStats (core_api_base.h:84-96,102-105) ---- */
B200_API int b200_set_target( b200_core* c, int width, int height, int spp );
B200_API int b200_set_point_light( b200_core* c, const float* position3, const float* radiance3 );
B200_API int b200_set_sky( b200_core* c, const float* radiance3 );
/* CoreAPI_Base::SetLights (core_api_base.h:102-105): the four arrays in the reference's layouts (common_classes.h:275-345:
 * CoreLightTri 96 B, CorePointLight 32 B, CoreSpotLight 48 B, CoreDirectionalLight 32 B); copied before returning */
B200_API int b200_set_lights( b200_core* c, const void* areaLights, int areaLightCount, const void* pointLights, int pointLightCount,
	const void* spotLights, int spotLightCount, const void* directionalLights, int directionalLightCount );
/* CoreAPI_Base::SetSkyData (core_api_base.h:106): equirectangular float3 pixels[width * height] (tools_shared.h:185-192) */
B200_API int b200_set_sky_data( b200_core* c, const float* rgb, int width, int height );
/* CoreAPI_Base::SetMaterials (core_api_base.h:100): CoreMaterial[count], 688 bytes each; the Lambert stage keeps color.value and
 * roughness.value (a colour component above 1 makes the material a light, common_classes.h ShadingData::IsEmissive) */
B200_API int b200_set_materials( b200_core* c, const void* coreMaterials, int count );
/* headless shortcut for hosts without CoreMaterial structs: (albedo rgb, roughness) per material index (CoreTri::material) */
B200_API int b200_set_material_colors( b200_core* c, const float* rgba, int count );
/* the full record the shade stage keeps, 16 floats per material: albedo rgb, roughness | transmission, eta, flags (1 SMOOTH +
 * 2 HASALPHA), 0 | extinction rgb (1 - CoreMaterial::absorption), diffuse texture index or -1 | uv scale xy, uv offset xy */
B200_API int b200_set_material_table( b200_core* c, const float* props16, int count );
/* CoreAPI_Base::SetTextures (core_api_base.h:98): CoreTexDesc[count] in the host layout (40 bytes, common_classes.h:246-269); call before
 * the materials.  ARGB32 / NRM32 / ARGB128 (HDR, float4 texels) maps are copied (MIP levels included) and sampled by the materials that name them */
/* Blue-noise sampling (BLUENOISE, rendercore_optix7/core_settings.h:29): the table the reference's cores build in Init
 * (rendercore_optix7/rendercore.cpp:258-267; 5 * 65536 words: Sobol sequence, scrambling tile at 65536, ranking tile at 3 * 65536).
 * With a table set, the first 256 samples of a pixel draw the pixel / lens position and the BSDF sample from it and the first two
 * the light sample (.optix.cu:66-79, pathtracer.h:168-222), as the reference does; words = NULL returns to xor32 numbers. */
B200_API int b200_set_blue_noise( b200_core* c, const uint32_t* words, int count );
/* Diagnostic: evaluate the Disney BSDF of the shade stage (Setting "bsdf" = 1) on its own, on the device.  in28: 28 floats per
 * item (color, transmittance, three 0.8 fixed-point parameter words as bits, eta, iN, N, iT, wow, then distance r0 r1 for mode 0 =
 * SampleBSDF or wiw for mode 1 = EvaluateBSDF; lib/sharedBSDFs/disney.h:227,314); out8: wiw, pdf, value, specular. */
B200_API int b200_disney_probe( b200_core* c, int mode, int n, const float* in28, float* out8 );
B200_API int b200_set_textures( b200_core* c, const void* coreTexDescs, int count );
/* multi-GPU: render into a caller-owned device accumulator float4[w*h] (e.g. a torch tensor that NCCL reduces);
 * NULL = the core's own.  After the gather, b200_finalize( total samples ) rebuilds the frame from it. */
B200_API int b200_set_accumulator_buffer( b200_core* c, void* dAccumulator );
B200_API int b200_finalize( b200_core* c, uint32_t totalSamples );
/* Multi-GPU gather over peer memory, for tile-sharded frames (Setting shardIndex / shardCount), one process per GPU of one box.
 * Once the peers are open (b200_open_peers) and the frame is sharded over exactly them, b200_render ENDS with the gather: the arrival
 * barrier, the pull + finalize kernel and the second barrier are the tail of the frame's enqueue (one submission, one synchronisation;
 * b200_stats.gatherMs / gatherKernelMs), so the frame and the accumulator a rank reads after b200_render are complete -- what a
 * RenderSystem calling CoreAPI::Render on every rank expects.  Setting gatherInRender = 0 leaves the gather to an explicit
 * b200_gather_finalize call.
 * b200_peer_handle (after b200_set_target, core-owned accumulator) returns 128 bytes: the CUDA IPC handles of this core's accumulator
 * and of its arrival flags.  The host exchanges the blobs between the ranks (any transport) and gives every core the whole table with
 * b200_open_peers (entry r = rank r's blob).  Per frame, after b200_render, b200_gather_finalize replaces all-reduce + b200_finalize:
 * ONE kernel pulls every foreign tile out of its owner's memory over NVLink, stores it into the local accumulator and writes the
 * finalized frame; the ranks meet on flags in peer memory before and after the pull (a missing peer is an error after ~2 s, not a hang).
 * The reference has no multi-GPU path (lib/CUDA/shared_host_code/cudatools.h:62-87 picks one device). */
B200_API int b200_peer_handle( b200_core* c, void* out128 );
B200_API int b200_open_peers( b200_core* c, const void* handles, int world, int rank );
B200_API int b200_close_peers( b200_core* c );
B200_API int b200_gather_finalize( b200_core* c, uint32_t totalSamples );
/* view: ViewPyramid (68 bytes, common_classes.h:362-385).  converge: 0 = Converge, 1 = Restart */
B200_API int b200_render( b200_core* c, const void* viewPyramid, int converge );
/* accumulator float4[w*h] (sum over samples; w = sample count) and the normalised RGBA32F frame */
B200_API int b200_read_accumulator( b200_core* c, float* dst );
B200_API int b200_read_frame( b200_core* c, float* dst );
/* the same without waiting: the frame leaves on a stream of its own (dst: page-locked host memory, or the copy is not asynchronous) while the next
 * b200_render runs; b200_wait_frame returns when the last frame asked for has arrived.  (The reference presents through a GL texture,
 * rendercore.cpp:857-870: a host that wants the pixels has no faster way there either.) */
B200_API int b200_read_frame_async( b200_core* c, float* dst );
B200_API int b200_wait_frame( b200_core* c );
B200_API void* b200_device_accumulator( b200_core* c ); /* device pointer, for NCCL gathers by the host layer */
/* stats: the CoreStats fields (core_api_base.h:30-61) flattened; see struct below */
typedef struct b200_stats {
	uint32_t smCount, ccMajor, ccMinor, vramMB;
	float bvhBuildTime;                 /* ms, host SAH build + conversion + upload (last SetGeometry batch) */
	float refitTime, tlasTime;          /* ms */
	uint32_t totalRays, totalExtensionRays, totalShadowRays;
	float renderTime;                   /* ms, whole Render() */
	uint32_t primaryRayCount; float traceTime0;
	uint32_t bounce1RayCount; float traceTime1;
	uint32_t deepRayCount; float traceTimeX;
	float shadowTraceTime, shadeTime;
	int32_t probedInstid, probedTriid; float probedDist;
	uint32_t wideNodeCount, triangleCount, instanceCount, tlasNodeCount;
	uint32_t kernelLaunches;            /* kernels this core has launched since creation */
	float lastTraceMs;                  /* device time of the last b200_trace_* kernel (CUDA events on the core's stream) */
	uint32_t bvhCacheHits;              /* meshes whose host-built tree came from the on-disk cache (environment B200_BVH_CACHE = directory) */
	float gatherKernelMs;               /* device time of the last gather's pull + finalize kernel alone (without the arrival waits) */
	uint32_t frameGraphLaunches;        /* frames replayed as a CUDA graph since creation (Setting frameGraph) */
	uint32_t frameChains, framePasses;  /* of the last b200_render: independent wave chains, passes the samples were split into */
	float hostEnqueueMs, hostWaitMs, hostTotalMs; /* host clock of the last b200_render: until everything was enqueued, waiting for the device, whole call */
	float gatherMs;                     /* of the last b200_render that ended with the peer-memory gather: device time from this rank's arrival to the end of the
	                                       second arrival (waiting for the slowest rank + pull + finalize); 0 when the frame had no gather */
	uint32_t argb32TexelCount, argb128TexelCount, nrm32TexelCount; /* CoreStats (core_api_base.h:38-40): texels of the last b200_set_textures per storage */
	float filterTime;                   /* ms, the reconstruction filter + TAA of the last b200_render (Setting filterCore); 0 when it did not run */
} b200_stats;
B200_API int b200_get_stats( b200_core* c, b200_stats* out );
B200_API int b200_set_probe( b200_core* c, int x, int y );

/* Filter mode.  Setting "filterCore" = 1 makes the core render as the reference's lib/RenderCore_Optix7Filter does: the keys "filter", "TAA",
 * "clampDirect", "clampIndirect" (sent to every core every frame, rendersystem.cpp:231-236) are then honoured (rendercore.cpp:648-670) -- paths
 * stop after one diffuse bounce, the accumulator restarts every frame, the first sample writes per-pixel features and the frame goes through
 * the reconstruction filter and TAA (filter.cu).  With filterCore = 0 (default) the four keys are accepted and ignored, as lib/rendercore_optix7
 * does.  b200_read_filter_buffer: the per-pixel buffers of the last filtered frame -- which = 0 features (uint4: albedo 10:11:11, packed normal,
 * depth bits, history | specular << 4 | material << 6), 1 world position + packed normal, 2 depth derivatives (zw), 3 indirect accumulator
 * plane, 4 motion vectors (float2), 5 luminance moments, 6 first reconstruction pass (packed), 7 TAA history; 16 bytes per pixel, motion 8. */
B200_API int b200_read_filter_buffer( b200_core* c, int which, void* dst );

/* ---- the SVGF-style filter and TAA, stage by stage, on caller-owned DEVICE buffers of w * h elements (current device, default stream,
 * synchronous).  b200_render runs the same kernels on the core's own buffers when Setting "filter" is on; these entry points are for hosts
 * that keep the buffers themselves and for the tests, which hold every stage against the reference's kernels.  Replaces prepareFilter /
 * applyFilter / TAApass / unsharpenTAA / finalizeNoTAA (lib/CUDA/shared_kernel_code/finalize_shared.h:178-196, 371-377, 425-431, 463-467,
 * 481-485).  Buffers: accumulator float4[2wh] (direct plane, indirect plane), features uint4, worldPos / deltaDepth / shading / moments /
 * A / B / C / pixels / frame float4, motion float2; prevView = pos, p1, p2, p3 of the previous frame's view pyramid (12 floats, host). */
B200_API int b200_filter_prepare( const void* accumulator, void* features, const void* worldPos, const void* prevWorldPos, void* shading, void* motion, void* moments,
	const void* prevMoments, const void* deltaDepth, const float* prevView, float j0, float j1, float prevj0, float prevj1, int w, int h, unsigned spp,
	float directClamp, float indirectClamp, int camIsStationary );
B200_API int b200_filter_reconstruct( const void* features, const void* prevWorldPos, const void* worldPos, const void* deltaDepth, const void* motion,
	const void* moments, const void* A, const void* B, void* C, int w, int h, int phase, int lastPass );
B200_API int b200_filter_taa( const void* pixels, const void* prevPixels, const void* motion, void* out, int w, int h ); /* out != pixels */
B200_API int b200_filter_unsharpen( const void* pixels, void* frame, int w, int h );
B200_API int b200_filter_finish_no_taa( const void* pixels, void* frame, int w, int h );

/* the core behind CreateCore() (NULL before CreateCore): lets a headless host read frames / stats of the plugin instance */
B200_API b200_core* b200_plugin_core( void );

/* pinned host memory helpers for callers that want the H2D/D2H legs at full PCIe speed */
B200_API void* b200_host_alloc( uint64_t bytes );
B200_API void b200_host_free( void* p );

/* ---- the plugin symbols themselves (declared for completeness; C++ linkage of the return type) ---- */
#ifdef __cplusplus
} /* extern "C" */
namespace lh2abi { class CoreAPI_Base; }
extern "C" B200_API lh2abi::CoreAPI_Base* CreateCore();
extern "C" B200_API void DestroyCore();
#endif

#endif /* RENDERCORE_B200_H */
