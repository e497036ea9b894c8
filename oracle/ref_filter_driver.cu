// oracle/ref_filter_driver.cu -- TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// C-ABI driver around the UNMODIFIED SVGF-style filter and TAA kernels of the reference, compiled IN PLACE (never copied) from
// /root/reference by oracle/Makefile with nvcc into oracle/_ref/liboracle_filter_ref.so:
//   prepareFilterKernel, applyFilterKernel, TAApassKernel, unsharpenTAAKernel, finalizeNoTAAKernel
//                                                               lib/CUDA/shared_kernel_code/finalize_shared.h:102-483
//   ReadWorldPos, ReadTexelConsistent(2), ReadTexelBmitchellNetravali   lib/CUDA/shared_kernel_code/sampling_shared.h:96-205
//   PackNormal2 / UnpackNormal2, RGBToYCoCg, HDRtoRGB32, CombineToFloat4 ...      lib/CUDA/shared_kernel_code/tools_shared.h:122-334
// They are the checker of lighthouse2_b200/csrc/filter.cu: the kernels are CUDA, so this checker runs on the GPU
// box (the .so travels there like the other files of oracle/_ref); tests/test_filter_gpu.py feeds both sides the same buffers.
// What the build needs beside the three headers is supplied here the way the cores' .cuda.cu / .cuda.h do it
// (lib/RenderCore_Optix7Filter/kernels/.cuda.cu:21-36, .cuda.h:19-57): the scalar typedefs, LH2_DEVFUNC, NEXTMULTIPLEOF, the
// __constant__ tables the sampling helpers name, and the render target.  The reference declares the target as a legacy surface
// REFERENCE (`surface<void, cudaSurfaceType2D>`, .cuda.h:52), which CUDA 12 no longer has; a surface OBJECT of the same name takes the
// same surf2Dwrite calls (the kernels are untouched), backed here by a plain cudaArray that the driver copies out.
#include <cuda_runtime.h>
#include <stdio.h>
typedef unsigned int uint;
typedef unsigned char uchar;
#include "helper_math.h"
#include "cuda_fp16.h"
#include "common_settings.h"
#include "common_classes.h"
#define LH2_DEVFUNC static __forceinline__ __device__
#define NEXTMULTIPLEOF( a, b ) (((a) + ((b) - 1)) & (0x7fffffff - ((b) - 1)))
__device__ cudaSurfaceObject_t renderTarget;

namespace lh2core {
__constant__ uchar4* argb32;
__constant__ float4* argb128;
__constant__ uchar4* nrm32;
__constant__ float3* skyPixels;
__constant__ int skywidth;
__constant__ int skyheight;
__constant__ float4* debugData;
#define BILINEAR
#include "tools_shared.h"
#include "sampling_shared.h"
#include "finalize_shared.h"
} // namespace lh2core

namespace {
cudaArray_t gTarget = 0;
cudaSurfaceObject_t gSurface = 0;
int gW = 0, gH = 0;
float4* gDebug = 0;
int ensureTarget( int w, int h )
{
	if (w == gW && h == gH) return 0;
	if (gSurface) cudaDestroySurfaceObject( gSurface ), cudaFreeArray( gTarget ), cudaFree( gDebug );
	cudaChannelFormatDesc desc = cudaCreateChannelDesc<float4>();
	if (cudaMallocArray( &gTarget, &desc, w, h, cudaArraySurfaceLoadStore ) != cudaSuccess) return 1;
	cudaResourceDesc res = {};
	res.resType = cudaResourceTypeArray, res.res.array.array = gTarget;
	if (cudaCreateSurfaceObject( &gSurface, &res ) != cudaSuccess) return 1;
	if (cudaMemcpyToSymbol( renderTarget, &gSurface, sizeof( gSurface ) ) != cudaSuccess) return 1;
	// prepareFilterKernel writes a debug pixel per thread (finalize_shared.h:125)
	if (cudaMalloc( &gDebug, (size_t)w * h * 16 ) != cudaSuccess) return 1;
	if (cudaMemcpyToSymbol( lh2core::debugData, &gDebug, sizeof( gDebug ) ) != cudaSuccess) return 1;
	gW = w, gH = h;
	return 0;
}
int finish() { return cudaDeviceSynchronize() == cudaSuccess && cudaGetLastError() == cudaSuccess ? 0 : 1; }
} // namespace

extern "C" {

// all pointers are DEVICE pointers; view = pos, p1, p2, p3 of the previous frame's pyramid (12 floats, host)
int ref_prepare_filter( const float4* accumulator, uint4* features, const float4* worldPos, const float4* prevWorldPos, float4* shading, float2* motion,
	float4* moments, float4* prevMoments, const float4* deltaDepth, const float* view, float j0, float j1, float prevj0, float prevj1, int w, int h,
	unsigned spp, float directClamp, float indirectClamp, int camIsStationary )
{
	if (ensureTarget( w, h )) return 1;
	ViewPyramid prev;
	prev.pos = make_float3( view[0], view[1], view[2] ), prev.p1 = make_float3( view[3], view[4], view[5] );
	prev.p2 = make_float3( view[6], view[7], view[8] ), prev.p3 = make_float3( view[9], view[10], view[11] );
	lh2core::prepareFilter( accumulator, features, worldPos, prevWorldPos, shading, motion, moments, prevMoments, deltaDepth, prev, j0, j1, prevj0, prevj1,
		w, h, spp, directClamp, indirectClamp, camIsStationary );
	return finish();
}

int ref_apply_filter( const uint4* features, const float4* prevWorldPos, const float4* worldPos, const float4* deltaDepth, const float2* motion,
	const float4* moments, const float4* A, const float4* B, float4* C, unsigned w, unsigned h, int phase, unsigned lastPass )
{
	lh2core::applyFilter( features, prevWorldPos, worldPos, deltaDepth, motion, moments, A, B, C, w, h, phase, lastPass );
	return finish();
}

int ref_taa_pass( float4* pixels, float4* prevPixels, const float4* worldPos, const float4* prevWorldPos, const float2* motion, unsigned w, unsigned h )
{
	lh2core::TAApass( pixels, prevPixels, 0, 0, worldPos, prevWorldPos, motion, w, h );
	return finish();
}

// the two kernels that end a filtered frame write the render target; `out` (device, w * h float4) receives it.  The target is cleared
// first: both kernels leave the one-pixel border untouched.
static int readTarget( float4* out, unsigned w, unsigned h )
{
	if (finish()) return 1;
	return cudaMemcpy2DFromArray( out, (size_t)w * 16, gTarget, 0, 0, (size_t)w * 16, h, cudaMemcpyDeviceToDevice ) == cudaSuccess ? 0 : 1;
}
static int clearTarget( unsigned w, unsigned h )
{
	if (ensureTarget( (int)w, (int)h )) return 1;
	float4* zero = 0;
	if (cudaMalloc( &zero, (size_t)w * h * 16 ) != cudaSuccess) return 1;
	cudaMemset( zero, 0, (size_t)w * h * 16 );
	const cudaError_t e = cudaMemcpy2DToArray( gTarget, 0, 0, zero, (size_t)w * 16, (size_t)w * 16, h, cudaMemcpyDeviceToDevice );
	cudaFree( zero );
	return e == cudaSuccess ? 0 : 1;
}
int ref_unsharpen_taa( const float4* pixels, float4* out, unsigned w, unsigned h )
{
	if (clearTarget( w, h )) return 1;
	lh2core::unsharpenTAA( pixels, w, h );
	return readTarget( out, w, h );
}
int ref_finalize_no_taa( float4* pixels, float4* out, unsigned w, unsigned h )
{
	if (clearTarget( w, h )) return 1;
	lh2core::finalizeNoTAA( pixels, w, h );
	return readTarget( out, w, h );
}

} // extern "C"
