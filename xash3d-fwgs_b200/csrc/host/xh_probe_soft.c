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
	if( height ) *height = (int)c->height;
	return c->data;
}
