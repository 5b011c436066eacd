/*
 * xh_probe_soft.c - ORACLE-ONLY probe: reads ref_soft's surface cache blocks.
 *
 * TEST INFRASTRUCTURE.  ref_soft keeps its cache pointers in
 * msurface_t.info->reserved[0..3] (CACHESPOT, ref/soft/r_local.h:968) and the
 * block header is surfcache_t (r_local.h:781-793).  We include the reference's
 * own header for the layout instead of restating it.  Only meaningful when the
 * loaded renderer is oracle/_ref/libref_soft.so.
 */
#define REF_DLL 1
#include "r_local.h"
#include "xh_host.h"

const void *xh_soft_cache_block( int surf, int mip, int *width, int *height )
{
	model_t *m = (model_t *)xh_world_model();
	if( !m || surf < 0 || surf >= m->numsurfaces || mip < 0 || mip > 3 )
		return NULL;
	surfcache_t *c = CACHESPOT(( &m->surfaces[surf] ))[mip];
	if( !c )
		return NULL;
	if( width ) *width = (int)c->width;
	/* surfcache_t.height is a debug field counted in bytes per row (r_surf.c:896); the block D_CacheSurface
	 * fills is lightextents[1] >> mip rows (r_surf.c:1126) */
	if( height ) *height = m->surfaces[surf].info->lightextents[1] >> mip;
	return c->data;
}

/*
 * ORACLE-ONLY driver loop: renders n camera views (x y z pitch yaw roll each) one after another through
 * the loaded ref_soft's GL_RenderFrame, the way the engine would, and copies vid.buffer / d_pzbuffer out
 * after each.  Both buffers are zeroed before every view when clear != 0 (ref_soft never clears,
 * r_edge.c:983) and the cached view cluster is forgotten so every view rebuilds its PVS marks
 * (R_MarkLeaves early-out, r_main.c:1023).  Returns the seconds spent inside GL_RenderFrame only.
 */
#include <time.h>

double xh_soft_render_views( const float *cams, int n, float fov_x, float fov_y, int x, int y, int w, int h, int flags, int clear,
	unsigned short *color_out, short *z_out )
{
	viddef_t *pvid = (viddef_t *)xh_ref_symbol( "vid" );
	short **pz = (short **)xh_ref_symbol( "d_pzbuffer" );
	int *oldcluster = (int *)xh_ref_symbol( "r_oldviewcluster" );
	if( !pvid || !pz || !oldcluster )
		return -1.0;
	size_t px = (size_t)pvid->width * pvid->height;
	double total = 0;
	for( int i = 0; i < n; i++ )
	{
		const float *c = cams + 6 * i;
		struct timespec t0, t1;
		if( clear )
		{
			memset( pvid->buffer, 0, px * sizeof( pixel_t ));
			memset( *pz, 0, px * sizeof( short ));
		}
		*oldcluster = -2;
		clock_gettime( CLOCK_MONOTONIC, &t0 );
		xh_render_view( c, c + 3, fov_x, fov_y, x, y, w, h, flags );
		clock_gettime( CLOCK_MONOTONIC, &t1 );
		total += ( t1.tv_sec - t0.tv_sec ) + 1e-9 * ( t1.tv_nsec - t0.tv_nsec );
		if( color_out ) memcpy( color_out + px * i, pvid->buffer, px * sizeof( pixel_t ));
		if( z_out ) memcpy( z_out + px * i, *pz, px * sizeof( short ));
	}
	return total;
}

/*
 * ORACLE-ONLY: the display tail of the reference.  Renders the views one after another like xh_soft_render_views and
 * ends the frame after each (R_EndFrame -> R_BlitScreen, ref/soft/r_main.c:1182, r_glblit.c:591), copying the
 * software surface (format: xh_set_sw_format) to display_out[n][h][w].
 */
int xh_soft_blit_views( const float *cams, int n, float fov_x, float fov_y, int x, int y, int w, int h, int flags, void *display_out )
{
	viddef_t *pvid = (viddef_t *)xh_ref_symbol( "vid" );
	short **pz = (short **)xh_ref_symbol( "d_pzbuffer" );
	int *oldcluster = (int *)xh_ref_symbol( "r_oldviewcluster" );
	int sw_w, sw_h, bpp;
	if( !pvid || !pz || !oldcluster )
		return -1;
	size_t px = (size_t)pvid->width * pvid->height;
	for( int i = 0; i < n; i++ )
	{
		const float *c = cams + 6 * i;
		memset( pvid->buffer, 0, px * sizeof( pixel_t ));
		memset( *pz, 0, px * sizeof( short ));
		*oldcluster = -2;
		xh_begin_frame();
		xh_render_view( c, c + 3, fov_x, fov_y, x, y, w, h, flags );
		xh_end_frame();
		const void *sw = xh_sw_buffer( &sw_w, &sw_h, &bpp );
		if( !sw || (size_t)sw_w * sw_h != px )
			return -2;
		memcpy((char *)display_out + px * bpp * i, sw, px * bpp );
	}
	return 0;
}
