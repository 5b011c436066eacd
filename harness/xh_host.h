/*
 * xh_host.h - headless engine stand-in ("host") for Xash3D FWGS renderer plugins.
 *
 * The real engine (engine/client/dll_int/ref_common.c:547-600) dlopen()s
 * libref_<name>.so, resolves GetRefAPI and hands it a ref_api_t of engine
 * callbacks.  The engine client cannot be built on this box (no SDL, missing
 * submodules), so this small C library plays the engine's part for any
 * renderer .so: the reference's own ref_soft (the parity oracle / CPU
 * baseline) and our ref_cuda.  It owns the model_t, the entity list, cvars,
 * gamma tables and light styles, exactly the things engine/ owns upstream.
 *
 * All exported functions are plain C (ctypes-friendly); state is per loaded
 * copy of this library (the renderer ABI has no context pointer, so one
 * library image == one engine instance; load a second copy for a second
 * renderer in the same process).
 */
#ifndef XH_HOST_H
#define XH_HOST_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define XH_EXPORT __attribute__((visibility("default")))

/* dlopen renderer, GetRefAPI handshake, R_Init, builtin textures. 0 on success. */
XH_EXPORT int xh_init( const char *ref_path, int width, int height );
XH_EXPORT void xh_shutdown( void );

/* cvars: must be set before xh_init for registration-time values, any time otherwise */
XH_EXPORT void xh_cvar_set( const char *name, const char *value );
XH_EXPORT float xh_cvar_get( const char *name );

/* load a BSP30 image from memory as the world (models[1]) + its inline submodels; calls R_NewMap */
XH_EXPORT int xh_load_map( const void *bsp, size_t len, const char *name );

/* time + light styles (engine/client/cl_tent.c:2390 CL_SetLightstyle, cl_frame.c:1300) */
XH_EXPORT void xh_set_time( double time, double oldtime );
XH_EXPORT void xh_set_lightstyle( int style, const char *pattern, float time );
XH_EXPORT void xh_run_lightstyles( void );

/* one view through ref.GL_RenderFrame (ref_viewpass_t, common/ref_params.h:97) */
XH_EXPORT void xh_begin_frame( void );
XH_EXPORT void xh_render_view( const float *origin, const float *angles, float fov_x, float fov_y,
	int x, int y, int w, int h, int flags );
XH_EXPORT void xh_end_frame( void );

/* entity list for the next views: R_ClearScene + R_AddEntity */
XH_EXPORT void xh_clear_scene( void );
XH_EXPORT int xh_add_brush_entity( int modelindex, const float *origin, const float *angles,
	int rendermode, int renderamt );

/* a one-frame sprite model from an RGBA image (type: SPR_* orientation, texformat: SPR_NORMAL / ADDITIVE / INDEXALPHA /
 * ALPHTEST, texflags: TF_* for the frame's texture); returns the model index */
XH_EXPORT int xh_create_sprite( const char *name, int w, int h, const void *rgba, int type, int texformat, int texflags );
/* a studio model from an MDL v10 file image; returns the model index */
XH_EXPORT int xh_load_studio( const char *name, const void *buffer, size_t len );
/* any model as an entity of the next views; ip[13]: modelindex, rendermode, renderamt, r, g, b, renderfx, frame, sequence,
 * body, skin, movetype, aiment; fp[9]: origin, angles, scale, framerate, animtime.  Returns the entity index */
XH_EXPORT int xh_add_entity( const int *ip, const float *fp );

/* R_Shutdown + R_Init on the loaded renderer (vid_restart); load the map again afterwards */
XH_EXPORT int xh_restart( void );

/* raw access for harnesses */
XH_EXPORT void *xh_ref_symbol( const char *name );      /* dlsym on the renderer .so */
XH_EXPORT void *xh_world_model( void );                  /* model_t* */
XH_EXPORT int xh_world_info( int *out, int n );          /* numsurfaces,numnodes,numleafs,numedges,numvertexes,... */
XH_EXPORT int xh_point_in_leaf( const float *p, int *contents, int *cluster );
XH_EXPORT int xh_surface_info( int surf, int *out, int n ); /* flags,extents,lightextents,styles,texnum,... */
XH_EXPORT const void *xh_gamma_table( int which, int *count ); /* 0 tex(u8) 1 light 2 linear 3 screen (u16) */
XH_EXPORT void xh_build_gamma( float gamma, float brightness, float texgamma, float lightgamma );
XH_EXPORT int xh_load_pic( const char *name, int w, int h, const void *rgba, int texflags );
XH_EXPORT int xh_frame_views_hud( const float *cams, int n, float fov_x, float fov_y, int x, int y, int w, int h, int flags,
	const float *hud, int nhud, void *display_out );
XH_EXPORT int xh_screenshot( int shot_type, void *out, size_t cap, int *width, int *height );
XH_EXPORT int xh_set_particles( const float *fp, int n );
/* the renderer's R_DecalShoot slot on the world model (entity 0, model 0): decal texture, centre, FDECAL_* flags, scale */
XH_EXPORT void xh_decal_shoot( int texture, const float *pos, int flags, float scale );
XH_EXPORT void xh_decal_shoot_entity( int texture, int entity, const float *pos, int flags, float scale );
XH_EXPORT int xh_decal_list( float *out, int max );
XH_EXPORT void xh_remove_decals( int texture );
XH_EXPORT void xh_clear_decals( void );
/* tracers (CL_DrawTracers): the sprite model of REF_DOT_SPRITE, the client's active tracers, the list after the renderer moved it */
XH_EXPORT int xh_world_to_screen( const float *p, float *out );
/* the wall EV_VisTraceLine finds (glow sprites): the plane coordinate[axis] = dist; axis < 0 removes it */
XH_EXPORT void xh_set_sight_wall( int axis, float dist );
XH_EXPORT void xh_set_dot_sprite( int modelindex );
XH_EXPORT int xh_set_tracers( const float *fp, int n );
XH_EXPORT int xh_get_tracers( float *out );
XH_EXPORT void xh_light_point( const float *p, int *out );
XH_EXPORT void xh_gamma_changed( float gamma, float brightness, float texgamma, float lightgamma );
XH_EXPORT uint64_t xh_callback_count( void );            /* engine callbacks hit since init (debug) */
XH_EXPORT const char *xh_last_error( void );
/* oracle-only probe: CACHESPOT(surf)[mip] of ref_soft -> surface cache block (r_local.h:781-793,968) */
XH_EXPORT const void *xh_soft_cache_block( int surf, int mip, int *width, int *height );
/* oracle-only driver loop: n cameras (x y z pitch yaw roll) through ref_soft's GL_RenderFrame, buffers copied out
 * per view; returns the seconds spent inside GL_RenderFrame (xh_probe_soft.c) */
XH_EXPORT double xh_soft_render_views( const float *cams, int n, float fov_x, float fov_y, int x, int y, int w, int h, int flags, int clear,
	unsigned short *color_out, short *z_out );
/* display format of the software surface handed out by SW_CreateBuffer (call before xh_init; default RGB565), and the
 * surface itself (what R_EndFrame wrote) */
XH_EXPORT void xh_set_sw_format( unsigned bpp, unsigned rmask, unsigned gmask, unsigned bmask );
XH_EXPORT const void *xh_sw_buffer( int *w, int *h, int *bpp );
/* entry i of the engine's dynamic light table (PARM_GET_DLIGHTS_PTR); radius 0 = off; alive while die >= time */
XH_EXPORT int xh_set_dlight( int i, const float *origin, float radius, int r, int g, int b, float die, float minlight );
/* the engine's frame loop, one view per frame (R_BeginFrame, GL_RenderFrame, R_EndFrame), any renderer plugin */
XH_EXPORT int xh_frame_views( const float *cams, int n, float fov_x, float fov_y, int x, int y, int w, int h, int flags, void *display_out );
/* oracle-only: the views one by one through GL_RenderFrame + R_EndFrame (R_BlitScreen); surface copied out per view */
XH_EXPORT int xh_soft_blit_views( const float *cams, int n, float fov_x, float fov_y, int x, int y, int w, int h, int flags, void *display_out );

#ifdef __cplusplus
}
#endif
#endif
