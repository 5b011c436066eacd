/*
 * ref_cuda.h - C ABI of libref_cuda.so, the B200-native renderer plugin for Xash3D FWGS.
 *
 * 1. The drop-in boundary is the reference's own plugin entry point: the engine dlopen()s
 *    "libref_cuda.so" for `-ref cuda` (engine/client/dll_int/ref_common.c:547-600), resolves
 *    GetRefAPI (engine/ref_api.h:660-661) and talks to the renderer only through the ref_interface_t
 *    table it fills (engine/ref_api.h:527-658).  Replaces ref/common/ref_context.c:57-81 +
 *    ref/soft/r_context.c:430-533.
 *
 * 2. Batch extension (additive; the engine never needs it): the stock ABI renders one view per
 *    GL_RenderFrame into one screen buffer.  Harnesses that want N independent observations per launch
 *    (RL-style rendering) resolve the RefCuda_* symbols with dlsym / COM_GetProcAddress.  Plain
 *    pointers and sizes only.
 */
#ifndef REF_CUDA_H
#define REF_CUDA_H

#ifdef __cplusplus
extern "C" {
#endif

struct ref_interface_s;
struct ref_api_s;
struct ref_globals_s;
struct ref_viewpass_s;

#define REF_CUDA_EXPORT __attribute__(( visibility( "default" )))

/* engine/ref_api.h:660 REFAPI.  Returns REF_API_VERSION (18) on success, 0 on version mismatch. */
REF_CUDA_EXPORT int GetRefAPI( int version, struct ref_interface_s *pFunctionTable, struct ref_api_s *engfuncs, struct ref_globals_s *pGlobals );

/*
 * Render nviews views (common/ref_params.h:97 ref_viewpass_t, the struct GL_RenderFrame takes) of the
 * current world in one batch.  Every view gets its own screen-sized (ref_globals_t width x height)
 * buffer: color_out[nviews][height][width] holds ref_soft pixel_t codes (ref/soft/r_local.h:84), z_out the
 * 16-bit 1/z buffer (d_pzbuffer); either may be NULL to leave the result in HBM.  clear != 0 starts from
 * zeroed buffers (ref_soft never clears: r_edge.c:983 D_SolidSurf skips the background surface).
 * Returns 0 on success; errors are also reported through the engine's Host_Error.
 */
REF_CUDA_EXPORT int RefCuda_RenderViews( const struct ref_viewpass_s *views, int nviews, unsigned short *color_out, short *z_out, int clear );

/*
 * The same batch, enqueued only: returns once the work is queued; color_out / z_out (pinned host memory, or the copies
 * degrade to synchronous ones) are complete after RefCuda_Wait.  A harness that renders batch after batch calls
 * RefCuda_RenderViewsAsync with alternating output buffers and RefCuda_Wait before it reads one: successive asynchronous
 * batches draw into alternating sets of device frame buffers, so the device->host copies of batch k (PCIe: the long pole of
 * the synchronous call) run under the stages of batch k + 1.  Every other RefCuda_* / ref_interface_t call waits first.
 * Batches with clear == 0, without host outputs or with fewer than 64 views run synchronously.
 */
REF_CUDA_EXPORT int RefCuda_RenderViewsAsync( const struct ref_viewpass_s *views, int nviews, unsigned short *color_out, short *z_out, int clear );
REF_CUDA_EXPORT int RefCuda_Wait( void );

/*
 * Display tail on the device (R_BuildScreenMap + R_BlitScreen, ref/soft/r_glblit.c:347-407,591-721): converts views
 * [first, first + nviews) of the last RefCuda_RenderViews batch, which stay in HBM, from pixel_t to a display format
 * described the way SW_CreateBuffer (engine/ref_api.h) describes the engine's surface - bpp 2 or 4 bytes and the three
 * channel masks - and writes them to dst[nviews * height][stride] (stride in pixels, >= width).  dst NULL: conversion
 * only.  ms_kernel (optional): CUDA-event milliseconds of the conversion kernel.  Returns 0 on success.
 */
REF_CUDA_EXPORT int RefCuda_BlitViews( int first, int nviews, void *dst, unsigned int stride, unsigned int bpp,
	unsigned int rmask, unsigned int gmask, unsigned int bmask, float *ms_kernel );

/*
 * No counterpart in the reference.  RefCuda_BlitViews' conversion kernel for
 * callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc): converts
 * npixels pixel_t codes (host memory) to the display format (bpp 2 or 4 and the three channel masks, as above) into dst
 * (host memory, npixels * bpp bytes).  Screen map, input and output are staged in a file-scope device array inside the
 * library (~12.7 MB per GPU that loads it, held until exit) - the call makes no cudaMalloc / cudaFree and no host
 * allocation.  Limit: npixels <= 2 073 600 (one 1920x1080 frame); more returns -2 with a message on the engine console.
 * The array is shared by all calls and keeps the last call's contents: calls that overlap on the GPU (several threads)
 * race; serialise them.  Works on the calling thread's current CUDA device.  Returns 0 on success.
 */
REF_CUDA_EXPORT int RefCuda_BlitPixelsNoAlloc( const unsigned short *pixels, unsigned int npixels, void *dst, unsigned int bpp,
	unsigned int rmask, unsigned int gmask, unsigned int bmask );

/* Re-run the device stages of the last batch with all inputs resident in HBM; CUDA-event milliseconds
 * per iteration: total and per stage, ms_stage[6] = [visibility+BSP walk, face/edge set-up, surface cache,
 * span generation, span drawing, entities (studio skinning + triangle rasteriser)].
 * flags: 1 = invalidate the surface cache before every iteration, 2 = flush L2 (256 MB write) between
 * iterations, outside the timed region, 8 = run the surface-cache kernels on the batch's stream instead of beside span
 * generation (times them alone; by default ms_stage[2] overlaps ms_stage[3]). */
REF_CUDA_EXPORT int RefCuda_BenchResident( int iters, int flags, float *ms_total, float *ms_stage );

/* Surface cache block of (surface, mip) as built on the device (D_CacheSurface, ref/soft/r_surf.c:1053). */
REF_CUDA_EXPORT int RefCuda_ReadSurfaceCache( int surf, int mip, unsigned short *out, int maxtexels, int *width, int *height, int *valid );

/* Counters of the last batch (rc_shared.h RC_CTR_*), kernels launched since R_Init, device name. */
REF_CUDA_EXPORT int RefCuda_GetCounters( unsigned long long *out, int n );
REF_CUDA_EXPORT unsigned long long RefCuda_LaunchCount( void );
/* host->device bytes the last RefCuda_RenderViews call uploaded: per-view constants, light style values, dynamic
 * lights, entity state (what a harness reports as the batch's input traffic) */
REF_CUDA_EXPORT unsigned long long RefCuda_BatchH2DBytes( void );
REF_CUDA_EXPORT const char *RefCuda_DeviceName( void );

#ifdef __cplusplus
}
#endif
#endif
