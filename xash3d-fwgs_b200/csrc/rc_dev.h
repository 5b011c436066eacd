/*
 * rc_dev.h - the thin C layer between the plugin's C host code (rc_host.c) and the sm_100a kernels
 * (rc_device.cu).  Plain pointers and sizes only.  Every function returns 0 on success, non-zero on a
 * CUDA error (text via rc_dev_last_error); there is no CPU implementation behind these entry points in
 * libref_cuda.so - without a CUDA device rc_dev_init fails and R_Init reports it.
 */
#ifndef RC_DEV_H
#define RC_DEV_H

#include <stddef.h>
#include "rc_shared.h"

#ifdef __cplusplus
extern "C" {
#endif

int rc_dev_init( int device );
void rc_dev_shutdown( void );
const char *rc_dev_last_error( void );
const char *rc_dev_name( void );

/* static tables: colormap[32*8192] u16, lightgamma[1024] u16, sintable[1280] i32 */
int rc_dev_upload_luts( const uint16_t *colormap, const uint16_t *lightgamma, const int32_t *sintable );
/* texture table + texel arena (pixel_t), replaces the previous one */
/* the texture table: tex[i] says where texture i's planes go in the pixel_t arena (ntexels long), src[i] where its RGBA8
 * source sits in rgba[] (nrgba uint32).  The arena is ENCODED ON THE DEVICE (k_tex_encode: GL_UploadTexture +
 * GL_BuildMipMap per output texel); textures with ofs[0] < 0 have no image */
int rc_dev_upload_textures( const rc_texture_t *tex, const rc_texsrc_t *src, int ntex, const uint32_t *rgba, size_t nrgba, size_t ntexels );
/* world arrays: sc holds HOST pointers and counts; cache_texels = size of the surface cache arena */
int rc_dev_upload_scene( const rc_scene_t *sc, size_t cache_texels );
/* texture number per surface after R_TextureAnimation (only changes when animated textures flip) */
int rc_dev_update_surface_texnums( const int32_t *texnum, int n );
/* blend tables of R_BuildBlendMaps (ref/soft/r_glblit.c:411-520): alphamap[3 * 1024 * 256] u16, addmap / modmap[65536] u8 */
int rc_dev_upload_blendluts( const uint16_t *alphamap, const uint8_t *addmap, const uint8_t *modmap );
/* the entities of the next batches as ordered draw items (rc_shared.h); nitems = 0: none */
/* the opaque brush entities of the next batches (R_DrawBEntitiesOnList); n = 0 clears them */
int rc_dev_set_bents( const rc_bent_t *bents, int n );
/* the HUD pass: the commands applied in order to frame `view` of the last batch (colour only) */
int rc_dev_draw2d( const rc_cmd2d_t *cmds, int n, int view );
/* the decals of the world's faces: records grouped by face in list order, per face {first record, count}, per face the
 * generation its cached blocks must carry (a face whose generation moved is rebuilt at every mip) */
int rc_dev_set_decals( const rc_ddecal_t *recs, int nrecs, const int32_t *face_first_count, const uint32_t *face_gen, int nfaces );
/* per-view records of the items with glow >= 0 (kRenderGlow sprites), laid out [view][nglow]; before the render call of those views */
int rc_dev_set_glow( const rc_glow_t *recs, int nglow, int nviews );
int rc_dev_set_items( const rc_item_t *items, int nitems, const rc_corner_t *corners, int ncorners, const rc_sortent_t *sortents, int nsortents );
/* studio models: the static arena (vertices, normals, bone indices, mesh corners of every submodel used so far; replaces the
 * previous one) and the per-batch state: drawn (entity, submodel) parts, bone matrices [n][12], bone-space light vectors [n][3] */
int rc_dev_upload_studio_static( const void *arena, size_t bytes );
int rc_dev_set_studio( const rc_spart_t *parts, int nparts, const float *bones, int nbones, const float *blightvecs, int totalverts, int totalnorms );
/* g_lightstylevalue[256] for the next batch (ref/common/ref_light.c:30-87) */
int rc_dev_set_lightstyles( const int32_t *values );
/* the active dynamic lights of the next batches (R_PushDlights, ref/common/ref_light.c:170: die >= time, radius != 0);
 * marks the world faces each one reaches (R_MarkLights) on the device.  n = 0: none */
int rc_dev_set_dlights( const rc_dlight_t *lights, int n );
/* drop every cached block (D_FlushCaches, ref/soft/r_surf.c:810) */
int rc_dev_flush_cache( void );

/*
 * Render nviews views of width x height (each view owns one width x height pixel_t buffer and z buffer).
 * color_out / z_out: HOST destinations [nviews][height][width] or NULL to leave results in HBM.
 * clear: non-zero -> framebuffers start as zeros (the reference never clears; harnesses zero both sides).
 */
int rc_dev_render( const rc_view_t *views, int nviews, int width, int height, int clear, uint16_t *color_out, int16_t *z_out );
/* the same, enqueued only: the outputs are complete after rc_dev_wait.  Successive asynchronous batches alternate between two
 * sets of frame buffers, so that the device->host copies of one batch run under the stages of the next */
int rc_dev_render_async( const rc_view_t *views, int nviews, int width, int height, int clear, uint16_t *color_out, int16_t *z_out );
int rc_dev_wait( void );

/* re-run the device stages of the last rc_dev_render batch (inputs resident in HBM) 'iters' times and
 * report CUDA-event milliseconds: total and per stage (vis+walk, faces, cache, spans, draw, entities).  The cache kernels run
 * on a side stream beside span generation: their time (own events) is not part of the sum of the other stages */
#define RC_BENCH_REBUILD_CACHE 1 /* invalidate every cache block before each iteration (config C3 behaviour) */
#define RC_BENCH_FLUSH_L2      2 /* write 256 MB between iterations, outside the timed region */
#define RC_BENCH_OVERLAP       4 /* run the batch as rc_dev_render does for host outputs: 4 view slices on 2 streams (total only) */
#define RC_BENCH_SERIAL_CACHE  8 /* surface-cache kernels on the batch's own stream instead of beside span generation: times them alone */
#define RC_BENCH_STAGES        6 /* vis+walk, faces, cache, spans, draw, entities */
int rc_dev_bench_resident( int iters, int flags, float *ms_total, float *ms_stage /* [RC_BENCH_STAGES] */ );

/* display tail (R_BuildScreenMap / R_BlitScreen, ref/soft/r_glblit.c:347,591): table[65536] maps pixel_t to the
 * display format (bpp 2: low 16 bits used; bpp 4).  rc_dev_blit converts views [first, first + nviews) of the last
 * batch, resident in HBM, and copies them to dst[nviews * height][stride] (pixels; NULL: conversion only);
 * ms_kernel (optional): CUDA-event time of the conversion kernel */
int rc_dev_set_screenmap( const uint32_t *table, int bpp );
int rc_dev_blit( int first, int nviews, void *dst, size_t stride, float *ms_kernel );
/* k_blit over a fixed file-scope device array instead of allocated buffers: table = the 65536-entry screen map, pixels =
 * npx pixel_t codes (host), dst = npx display pixels of bpp bytes (host).  Allocates nothing; npx <= RC_STATIC_BLIT_MAXPX. */
#define RC_STATIC_BLIT_MAXPX ( 1920 * 1080 )
int rc_dev_blit_static( const uint32_t *table, int bpp, const uint16_t *pixels, size_t npx, void *dst );

/* probes for the parity tests */
int rc_dev_read_cache_block( int surf, int mip, uint16_t *out, int maxtexels, int *width, int *height, int *valid );
int rc_dev_read_counters( uint64_t *out, int n );
int rc_dev_read_view_surfs( int view, rc_surf_t *out, int max, int *n );
int rc_dev_read_view_edges( int view, rc_edge_t *out, int max, int *n );
/* host->device bytes uploaded since the last call with reset != 0 */
uint64_t rc_dev_h2d_bytes( int reset );
/* launch counter: kernels of ours launched since init */
uint64_t rc_dev_launch_count( void );

#ifdef __cplusplus
}
#endif
#endif
