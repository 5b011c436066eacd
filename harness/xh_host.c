/*
 * xh_host.c - headless engine stand-in: ref_api_t callbacks + renderer loader.
 *
 * Plays the role of engine/client/dll_int/ref_common.c (R_LoadProgs :547,
 * R_Init :742, R_CreateBuiltinTextures :494) and the handful of engine
 * services a renderer consumes.  It is written from the
 * ABI in engine/ref_api.h; it is not a copy of engine code.
 */
#define _GNU_SOURCE
#include <dlfcn.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "xh_internal.h"
#include "pm_defs.h"
#include "sprite.h"
#include "studio.h"
#include "r_studioint.h"

xh_state_t xh;

/* ------------------------------------------------------------------ */
/* errors / console                                                     */

static char xh_errbuf[1024];

const char *xh_last_error( void ) { return xh_errbuf; }

static void xh_Con_Printf( const char *fmt, ... )
{
	xh.callbacks++;
	if( !xh.verbose ) return;
	va_list ap; va_start( ap, fmt ); vfprintf( stderr, fmt, ap ); va_end( ap );
}

static void xh_Con_Quiet( const char *fmt, ... ) { xh.callbacks++; (void)fmt; }

static void xh_Host_Error( const char *fmt, ... )
{
	va_list ap; va_start( ap, fmt ); vsnprintf( xh_errbuf, sizeof( xh_errbuf ), fmt, ap ); va_end( ap );
	fprintf( stderr, "[xh_host] Host_Error: %s\n", xh_errbuf );
	/* the engine longjmps to the console here; a test harness has no console to go to */
	abort();
}

/* ------------------------------------------------------------------ */
/* memory pools: plain malloc with a size header (for realloc-with-clear) */

typedef struct { size_t size; size_t pad; } xh_memhdr_t;

static poolhandle_t xh_Mem_AllocPool( const char *name, unsigned int flags, const char *file, int line )
{
	(void)name; (void)flags; (void)file; (void)line;
	return ++xh.npools;
}

static void xh_Mem_FreePool( poolhandle_t *pool, const char *file, int line )
{
	(void)file; (void)line;
	if( pool ) *pool = 0; /* allocations are leaked on purpose: process-lifetime harness */
}

static void *xh_Mem_Alloc( poolhandle_t pool, size_t size, qboolean clear, const char *file, int line )
{
	(void)pool; (void)file; (void)line;
	xh_memhdr_t *h = clear ? calloc( 1, sizeof( *h ) + size + 64 ) : malloc( sizeof( *h ) + size + 64 );
	if( !h ) xh_Host_Error( "out of memory (%zu bytes)", size );
	h->size = size;
	return h + 1;
}

static void xh_Mem_Free( void *data, const char *file, int line )
{
	(void)file; (void)line;
	if( data ) free((xh_memhdr_t *)data - 1 );
}

static void *xh_Mem_Realloc( poolhandle_t pool, void *ptr, size_t size, qboolean clear, const char *file, int line )
{
	if( !ptr ) return xh_Mem_Alloc( pool, size, clear, file, line );
	xh_memhdr_t *h = (xh_memhdr_t *)ptr - 1;
	size_t old = h->size;
	h = realloc( h, sizeof( *h ) + size + 64 );
	if( !h ) xh_Host_Error( "out of memory (%zu bytes)", size );
	if( clear && size > old ) memset((byte *)( h + 1 ) + old, 0, size - old );
	h->size = size;
	return h + 1;
}

void *xh_alloc( size_t size ) { return xh_Mem_Alloc( 0, size, true, __FILE__, __LINE__ ); }

/* ------------------------------------------------------------------ */
/* cvars: linked list of convar_t (common/cvardef.h:97)                 */

static convar_t *xh_cvar_find( const char *name )
{
	for( convar_t *c = xh.cvars; c; c = c->next )
		if( !strcasecmp( c->name, name )) return c;
	return NULL;
}

static void xh_cvar_assign( convar_t *c, const char *value )
{
	c->string = strdup( value );
	c->value = (float)atof( value );
	c->flags |= FCVAR_CHANGED;
}

static convar_t *xh_cvar_create( const char *name, const char *value, int flags )
{
	convar_t *c = xh_cvar_find( name );
	if( c ) return c;
	c = calloc( 1, sizeof( *c ));
	c->name = strdup( name );
	c->desc = (char *)"";
	c->def_string = strdup( value );
	c->flags = flags | FCVAR_ALLOCATED;
	xh_cvar_assign( c, value );
	c->flags &= ~FCVAR_CHANGED;
	c->next = xh.cvars; xh.cvars = c;
	return c;
}

/* a renderer-owned static convar_t (CVAR_DEFINE) being linked in (engine/common/cvar.c Cvar_RegisterVariable) */
static void xh_Cvar_RegisterVariable( convar_t *var )
{
	xh.callbacks++;
	convar_t *old = xh_cvar_find( var->name );
	const char *value = var->string;
	char *preset = NULL;
	if( old )
	{
		/* value preset by the harness before the renderer registered its own variable: keep it, unlink ours */
		preset = strdup( old->string );
		value = preset;
		for( convar_t **pp = &xh.cvars; *pp; pp = &( *pp )->next )
			if( *pp == old ) { *pp = old->next; break; }
	}
	var->def_string = var->string;
	var->string = strdup( value );
	var->value = (float)atof( value );
	var->next = xh.cvars; xh.cvars = var;
	free( preset );
}

static cvar_t *xh_Cvar_Get( const char *name, const char *value, int flags, const char *desc )
{
	(void)desc; xh.callbacks++;
	return (cvar_t *)xh_cvar_create( name, value, flags );
}

static cvar_t *xh_GetCvarPointer( const char *name ) { xh.callbacks++; return (cvar_t *)xh_cvar_find( name ); }
static float xh_GetCvarFloat( const char *name ) { convar_t *c = xh_cvar_find( name ); return c ? c->value : 0.0f; }
static const char *xh_GetCvarString( const char *name ) { convar_t *c = xh_cvar_find( name ); return c ? c->string : ""; }
static void xh_Cvar_Set( const char *name, const char *value )
{
	convar_t *c = xh_cvar_find( name );
	if( !c ) c = xh_cvar_create( name, value, 0 );
	xh_cvar_assign( c, value );
}
static void xh_Cvar_SetValue( const char *name, float value )
{
	char buf[64]; snprintf( buf, sizeof( buf ), "%g", value ); xh_Cvar_Set( name, buf );
}
static void xh_Cvar_FullSet( const char *name, const char *value, int flags ) { (void)flags; xh_Cvar_Set( name, value ); }

void xh_cvar_set( const char *name, const char *value ) { xh_Cvar_Set( name, value ); }
float xh_cvar_get( const char *name ) { return xh_GetCvarFloat( name ); }

/* ------------------------------------------------------------------ */
/* small engine services                                                */

static int xh_Cmd_AddCommand( const char *n, void (*f)( void ), const char *d ) { (void)n; (void)f; (void)d; return 1; }
static void xh_Cmd_RemoveCommand( const char *n ) { (void)n; }
static int xh_Cmd_Argc( void ) { return 0; }
static const char *xh_Cmd_Argv( int i ) { (void)i; return ""; }
static const char *xh_Cmd_Args( void ) { return ""; }
static void xh_Cbuf_Text( const char *t ) { (void)t; }
static void xh_Cbuf_Execute( void ) {}
static void xh_void( void ) { xh.callbacks++; }
static int xh_Sys_CheckParm( const char *flag ) { (void)flag; return 0; }
static qboolean xh_R_Init_Video( ref_graphic_apis_t type ) { return type == REF_SOFTWARE; }

/* deterministic LCG in place of engine/common/common.c COM_RandomLong: the value only feeds
 * rtable[][] (random tiling textures "-0name"), identical for every renderer under this host */
static void xh_SetRandomSeed( int seed ) { xh.rng = (uint32_t)seed; }
static int xh_RandomLong( int lo, int hi )
{
	xh.rng = xh.rng * 1664525u + 1013904223u;
	uint32_t span = (uint32_t)( hi - lo + 1 );
	return span ? lo + (int)(( xh.rng >> 8 ) % span ) : lo;
}
static float xh_RandomFloat( float lo, float hi )
{
	xh.rng = xh.rng * 1664525u + 1013904223u;
	return lo + ( hi - lo ) * (float)(( xh.rng >> 8 ) & 0xFFFF ) / 65535.0f;
}

static intptr_t xh_EngineGetParm( int parm, int arg )
{
	(void)arg; xh.callbacks++;
	switch( parm )
	{
	case PARM_GET_CLIENT_PTR: return (intptr_t)&xh.cl;
	case PARM_GET_HOST_PTR: return (intptr_t)&xh.host;
	case PARM_GET_MOVEVARS_PTR: return (intptr_t)&xh.movevars;
	case PARM_GET_PALETTE_PTR: return (intptr_t)xh.palette;
	case PARM_GET_VIEWENT_PTR: return (intptr_t)&xh.viewent;
	case PARM_GET_TEXGAMMATABLE_PTR: return (intptr_t)xh.texgamma;
	case PARM_GET_LIGHTGAMMATABLE_PTR: return (intptr_t)xh.lightgamma;
	case PARM_GET_SCREENGAMMATABLE_PTR: return (intptr_t)xh.screengamma;
	case PARM_GET_LINEARGAMMATABLE_PTR: return (intptr_t)xh.lineargamma;
	case PARM_GET_LIGHTSTYLES_PTR: return (intptr_t)xh.lightstyles;
	case PARM_GET_DLIGHTS_PTR: return (intptr_t)xh.dlights;
	case PARM_GET_ELIGHTS_PTR: return (intptr_t)xh.elights;
	case PARM_GET_WORLD_PTR: return 0;
	case PARM_CONNSTATE: return 5; /* ca_active, engine/client/client.h */
	case PARM_SINGLEPLAYER_GAME: return 1;
	case PARM_NUMENTITIES: return xh.num_entities;
	case PARM_THIRDPERSON: case PARM_DEV_OVERVIEW: case PARM_QUAKE_COMPATIBLE:
	case PARM_PLAYING_DEMO: case PARM_WATER_LEVEL: return 0;
	case PARM_LOCAL_HEALTH: return 100;
	default: return 0;
	}
}

/* software "window": the engine hands out an RGB565 surface (engine/platform/sdl2/vid_sdl2.c:52-229) */
static qboolean xh_SW_CreateBuffer( int w, int h, uint *stride, uint *bpp, uint *r, uint *g, uint *b )
{
	if( !xh.sw_bpp ) { xh.sw_bpp = 2; xh.sw_mask[0] = 0xF800; xh.sw_mask[1] = 0x07E0; xh.sw_mask[2] = 0x001F; }
	free( xh.swbuffer );
	xh.swbuffer = calloc((size_t)w * h, xh.sw_bpp );
	xh.sw_w = w; xh.sw_h = h;
	*stride = w; *bpp = xh.sw_bpp; *r = xh.sw_mask[0]; *g = xh.sw_mask[1]; *b = xh.sw_mask[2];
	return xh.swbuffer != NULL;
}
static void *xh_SW_LockBuffer( void ) { return xh.swbuffer; }

/* display format of the software surface handed to the renderer; call before xh_init (default RGB565) */
void xh_set_sw_format( unsigned bpp, unsigned rmask, unsigned gmask, unsigned bmask )
{
	xh.sw_bpp = bpp; xh.sw_mask[0] = rmask; xh.sw_mask[1] = gmask; xh.sw_mask[2] = bmask;
}

/* the software surface (what R_EndFrame / R_BlitScreen wrote) */
const void *xh_sw_buffer( int *w, int *h, int *bpp )
{
	if( w ) *w = xh.sw_w;
	if( h ) *h = xh.sw_h;
	if( bpp ) *bpp = (int)xh.sw_bpp;
	return xh.swbuffer;
}
static void xh_SW_UnlockBuffer( void ) {}

static void xh_R_StoreEfrags( struct efrag_s **pp, int framecount ) { (void)pp; (void)framecount; }
/* the client's effects pass (CL_DrawEFX, engine/client/cl_efx.c:2079): the particles the test placed go to the renderer in the
 * translucent pass, as the engine hands over cl_active_particles */
static particle_t xh_particles[4096];
static int xh_nparticles;
static void xh_CL_ThinkParticle( double frametime, particle_t *p ) { (void)frametime; (void)p; }
/* tracers: the list the test set and the copy handed to the renderer, which moves them (CL_DrawTracers).  Every view is a
 * frame of the same client state here, so each call starts from the list as set; xh_get_tracers reads the moved copy. */
static particle_t xh_tracers_set[1024], xh_tracers[1024];
static int xh_ntracers;
static model_t *xh_dot_sprite;
static model_t *xh_GetDefaultSprite( enum ref_defaultsprite_e spr ) { return spr == REF_DOT_SPRITE ? xh_dot_sprite : NULL; }
static void xh_CL_DrawEFX( float time, qboolean trans )
{
	xh.callbacks++;
	if( trans && xh_nparticles && xh.ref.CL_DrawParticles )
		xh.ref.CL_DrawParticles( time, xh_particles, 2.0f ); /* PART_SIZE = max( 0.5, cl_draw_particles ): 2 keeps them visible at 320x240 */
	if( trans && xh_ntracers && xh.ref.CL_DrawTracers )
	{
		memcpy( xh_tracers, xh_tracers_set, sizeof( particle_t ) * xh_ntracers );
		for( int i = 0; i < xh_ntracers; i++ )
			xh_tracers[i].next = i + 1 < xh_ntracers ? &xh_tracers[i + 1] : NULL;
		xh.ref.CL_DrawTracers( time, xh_tracers );
	}
}

/* the renderer's WorldToScreen slot (the view rendered last): out = {x, y, z}, returns its "behind the eye" flag */
int xh_world_to_screen( const float *p, float *out )
{
	vec3_t w = { p[0], p[1], p[2] }, s = { 0, 0, 0 };
	const int behind = xh.ref.WorldToScreen( w, s );
	out[0] = s[0]; out[1] = s[1]; out[2] = s[2];
	return behind;
}

/* EV_VisTraceLine for the glow sprites' sight line (engine/client/cl_game.c): this harness has no collision hulls - the
 * line is stopped by one wall the test places, the plane coordinate[axis] = dist (axis < 0: nothing in the way) */
static int xh_wall_axis = -1;
static float xh_wall_dist;
static pmtrace_t xh_vistrace;
static pmtrace_t *xh_EV_VisTraceLine( float *start, float *end, int flags )
{
	(void)flags;
	xh.callbacks++;
	memset( &xh_vistrace, 0, sizeof( xh_vistrace ));
	xh_vistrace.fraction = 1.0f;
	if( xh_wall_axis >= 0 && xh_wall_axis < 3 )
	{
		const float a = start[xh_wall_axis] - xh_wall_dist, b = end[xh_wall_axis] - xh_wall_dist;
		if(( a < 0 ) != ( b < 0 ) && a != b )
			xh_vistrace.fraction = a / ( a - b );
	}
	for( int i = 0; i < 3; i++ )
		xh_vistrace.endpos[i] = start[i] + xh_vistrace.fraction * ( end[i] - start[i] );
	return &xh_vistrace;
}

void xh_set_sight_wall( int axis, float dist )
{
	xh_wall_axis = axis; xh_wall_dist = dist;
}

/* the model GetDefaultSprite( REF_DOT_SPRITE ) answers with (the engine's sprites/dot.spr) */
void xh_set_dot_sprite( int modelindex )
{
	xh_dot_sprite = ( modelindex > 0 && modelindex <= xh.cl.nummodels ) ? xh.cl.models[modelindex] : NULL;
}

/* fp: n x {org[3], vel[3], ramp, die, colour index, type, alpha (particle_t::unused)} */
int xh_set_tracers( const float *fp, int n )
{
	if( n > 1024 ) n = 1024;
	memset( xh_tracers_set, 0, sizeof( xh_tracers_set ));
	for( int i = 0; i < n; i++ )
	{
		particle_t *p = &xh_tracers_set[i];
		const float *f = fp + 11 * i;
		p->org[0] = f[0]; p->org[1] = f[1]; p->org[2] = f[2];
		p->vel[0] = f[3]; p->vel[1] = f[4]; p->vel[2] = f[5];
		p->ramp = f[6]; p->die = f[7]; p->color = (short)f[8]; p->type = (ptype_t)(int)f[9]; p->unused = (short)f[10];
	}
	xh_ntracers = n;
	return n;
}

/* the list as the renderer's last CL_DrawTracers call left it: n x {org[3], vel[3], alpha} */
int xh_get_tracers( float *out )
{
	for( int i = 0; i < xh_ntracers; i++ )
	{
		const particle_t *p = &xh_tracers[i];
		float *f = out + 7 * i;
		f[0] = p->org[0]; f[1] = p->org[1]; f[2] = p->org[2]; f[3] = p->vel[0]; f[4] = p->vel[1]; f[5] = p->vel[2]; f[6] = (float)p->unused;
	}
	return xh_ntracers;
}

/* fp: n x {x, y, z, colour index, die time, type} */
void xh_decal_shoot( int texture, const float *pos, int flags, float scale )
{
	vec3_t p = { pos[0], pos[1], pos[2] };
	xh.ref.R_DecalShoot( texture, 0, 0, p, flags, scale );
}

/* the same on a brush entity: world-space point, the entity's number (its model is looked up by the renderer) */
void xh_decal_shoot_entity( int texture, int entity, const float *pos, int flags, float scale )
{
	vec3_t p = { pos[0], pos[1], pos[2] };
	xh.ref.R_DecalShoot( texture, entity, 0, p, flags, scale );
}

/* the renderer's R_CreateDecalList slot: n x {position[3], entityIndex, depth, flags, scale, impactPlaneNormal[3], name checksum} */
int xh_decal_list( float *out, int max )
{
	static decallist_t list[MAX_RENDER_DECALS];
	memset( list, 0, sizeof( list ));
	int n = xh.ref.R_CreateDecalList( list );
	if( n > max ) n = max;
	for( int i = 0; i < n; i++ )
	{
		float *f = out + 11 * i;
		const decallist_t *e = &list[i];
		unsigned h = 0;
		for( const char *c = e->name; *c; c++ ) h = h * 31u + (unsigned char)*c;
		f[0] = e->position[0]; f[1] = e->position[1]; f[2] = e->position[2]; f[3] = e->entityIndex; f[4] = e->depth; f[5] = e->flags;
		f[6] = e->scale; f[7] = e->impactPlaneNormal[0]; f[8] = e->impactPlaneNormal[1]; f[9] = e->impactPlaneNormal[2]; f[10] = (float)( h & 0xFFFFFu );
	}
	return n;
}

void xh_remove_decals( int texture )
{
	xh.ref.R_DecalRemoveAll( texture );
}

void xh_clear_decals( void )
{
	xh.ref.R_ClearAllDecals();
}

int xh_set_particles( const float *fp, int n )
{
	if( n > 4096 ) n = 4096;
	memset( xh_particles, 0, sizeof( xh_particles ));
	for( int i = 0; i < n; i++ )
	{
		particle_t *p = &xh_particles[i];
		p->org[0] = fp[6 * i]; p->org[1] = fp[6 * i + 1]; p->org[2] = fp[6 * i + 2];
		p->color = (short)fp[6 * i + 3]; p->die = fp[6 * i + 4]; p->type = (ptype_t)(int)fp[6 * i + 5];
		p->next = i + 1 < n ? &xh_particles[i + 1] : NULL;
	}
	xh_nparticles = n;
	return n;
}
static struct cl_entity_s *xh_R_BeamGetEntity( int i ) { (void)i; return NULL; }
/* engine/common/model.c Mod_Extradata: the cache data of a model of the asked-for type only */
static void *xh_Mod_Extradata( int type, model_t *m ) { return ( m && (int)m->type == type ) ? m->cache.data : NULL; }
static int xh_GL_SetAttribute( int a, int v ) { (void)a; (void)v; return 0; }
static int xh_GL_GetAttribute( int a, int *v ) { (void)a; if( v ) *v = 0; return 0; }
static void *xh_GL_GetProcAddress( const char *n ) { (void)n; return NULL; }
static double xh_Time( void ) { return xh.host.realtime; }
static void xh_Image_SetForceFlags( uint f ) { (void)f; }
static void xh_Image_ClearForceFlags( void ) {}
static qboolean xh_Image_CustomPalette( void ) { return false; }
byte *xh_mdl_pointer;
static void xh_Image_SetMDLPointer( byte *p ) { xh_mdl_pointer = p; }
static qboolean xh_Image_Process( rgbdata_t **pix, int w, int h, uint flags, float r ) { (void)pix; (void)w; (void)h; (void)flags; (void)r; return true; }

/* world as physent 0 (pmove->physents[0], engine/client/cl_pmove.c): light tracing walks it */
static struct physent_s *xh_EV_GetPhysent( int idx )
{
	static physent_t world;
	if( idx != 0 || !xh.world ) return NULL;
	memset( &world, 0, sizeof( world ));
	strcpy( world.name, "world" );
	world.model = xh.world;
	world.solid = SOLID_BSP;
	return &world;
}
static void xh_SetLocalLightLevel( int level ) { (void)level; }
static void *xh_null_ptr( void ) { return NULL; }

/* ------------------------------------------------------------------ */
/* trap stubs for every slot we do not implement: a call is a harness bug, say which */

#define XH_NSTUBS 128
static void xh_trap( int idx )
{
	fprintf( stderr, "[xh_host] renderer called unimplemented ref_api_t slot #%d (see engine/ref_api.h:368-520)\n", idx );
	abort();
}
#define S1( n ) static intptr_t xh_stub_##n( void ) { xh_trap( n ); return 0; }
#define S10( d ) S1( d##0 ) S1( d##1 ) S1( d##2 ) S1( d##3 ) S1( d##4 ) S1( d##5 ) S1( d##6 ) S1( d##7 ) S1( d##8 ) S1( d##9 )
S1( 0 ) S1( 1 ) S1( 2 ) S1( 3 ) S1( 4 ) S1( 5 ) S1( 6 ) S1( 7 ) S1( 8 ) S1( 9 )
S10( 1 ) S10( 2 ) S10( 3 ) S10( 4 ) S10( 5 ) S10( 6 ) S10( 7 ) S10( 8 ) S10( 9 ) S10( 10 ) S10( 11 )
S1( 120 ) S1( 121 ) S1( 122 ) S1( 123 ) S1( 124 ) S1( 125 ) S1( 126 ) S1( 127 )
#undef S1
#define S1( n ) xh_stub_##n,
static intptr_t (*const xh_stubs[XH_NSTUBS])( void ) = {
	S1( 0 ) S1( 1 ) S1( 2 ) S1( 3 ) S1( 4 ) S1( 5 ) S1( 6 ) S1( 7 ) S1( 8 ) S1( 9 )
	S10( 1 ) S10( 2 ) S10( 3 ) S10( 4 ) S10( 5 ) S10( 6 ) S10( 7 ) S10( 8 ) S10( 9 ) S10( 10 ) S10( 11 )
	S1( 120 ) S1( 121 ) S1( 122 ) S1( 123 ) S1( 124 ) S1( 125 ) S1( 126 ) S1( 127 )
};

static mspriteframe_t *xh_R_GetSpriteFrame( const model_t *pModel, int frame, float yaw );
static void *xh_R_StudioGetAnim( studiohdr_t *hdr, model_t *mod, mstudioseqdesc_t *pseqdesc );

/* screenshots: FS_SaveImage keeps the image the renderer hands over (VID_ScreenShot, ref/soft/r_glblit.c:748) */
static struct { rgbdata_t *pic; byte *data; int width, height, type; size_t size; } xh_shot;
static qboolean xh_FS_SaveImage( const char *filename, rgbdata_t *pix )
{
	(void)filename;
	free( xh_shot.data );
	xh_shot.pic = pix; xh_shot.width = pix->width; xh_shot.height = pix->height; xh_shot.type = pix->type; xh_shot.size = pix->size;
	xh_shot.data = malloc( pix->size );
	memcpy( xh_shot.data, pix->buffer, pix->size );
	return true;
}
static void xh_FS_FreeImage_any( rgbdata_t *pic )
{
	if( pic && pic == xh_shot.pic ) /* allocated by the renderer from a memory pool, not by the harness's loader */
	{
		xh_Mem_Free( pic->buffer, __FILE__, __LINE__ ); xh_Mem_Free( pic, __FILE__, __LINE__ );
		xh_shot.pic = NULL;
		return;
	}
	xh_FS_FreeImage( pic );
}
static const bpc_desc_t *xh_Image_GetPFDesc( int idx )
{
	static bpc_desc_t d;
	d.format = idx; d.bpp = ( idx == PF_RGB_24 || idx == PF_BGR_24 ) ? 3 : ( idx == PF_INDEXED_24 || idx == PF_INDEXED_32 ) ? 1 : 4;
	return &d;
}
static void xh_AllowDirectPaths( qboolean enable ) { (void)enable; }
static fs_api_t xh_fsapi;

/* VID_ScreenShot of the last frame; the image FS_SaveImage received -> out (BGRA), returns its byte size or < 0 */
int xh_screenshot( int shot_type, void *out, size_t cap, int *width, int *height )
{
	if( !xh.ref.VID_ScreenShot( "shot.bmp", shot_type ) || !xh_shot.data )
		return -1;
	if( width ) *width = xh_shot.width;
	if( height ) *height = xh_shot.height;
	if( xh_shot.size > cap ) return -2;
	memcpy( out, xh_shot.data, xh_shot.size );
	return (int)xh_shot.size;
}

static void xh_fill_engfuncs( ref_api_t *e )
{
	/* every slot is a pointer: point all of them at numbered traps first */
	void **slots = (void **)e;
	size_t n = sizeof( *e ) / sizeof( void * );
	for( size_t i = 0; i < n; i++ )
		slots[i] = (void *)xh_stubs[i < XH_NSTUBS ? i : XH_NSTUBS - 1];

	e->EngineGetParm = xh_EngineGetParm;
	e->Cvar_Get = xh_Cvar_Get;
	e->pfnGetCvarPointer = xh_GetCvarPointer;
	e->pfnGetCvarFloat = xh_GetCvarFloat;
	e->pfnGetCvarString = xh_GetCvarString;
	e->Cvar_SetValue = xh_Cvar_SetValue;
	e->Cvar_Set = xh_Cvar_Set;
	e->Cvar_RegisterVariable = xh_Cvar_RegisterVariable;
	e->Cvar_FullSet = xh_Cvar_FullSet;
	e->Cmd_AddCommand = xh_Cmd_AddCommand;
	e->Cmd_RemoveCommand = xh_Cmd_RemoveCommand;
	e->Cmd_Argc = xh_Cmd_Argc;
	e->Cmd_Argv = xh_Cmd_Argv;
	e->Cmd_Args = xh_Cmd_Args;
	e->Cbuf_AddText = xh_Cbuf_Text;
	e->Cbuf_InsertText = xh_Cbuf_Text;
	e->Cbuf_Execute = xh_Cbuf_Execute;
	e->Con_Printf = xh_Con_Printf;
	e->Con_DPrintf = xh_Con_Quiet;
	e->Con_Reportf = xh_Con_Quiet;
	e->R_BeamGetEntity = xh_R_BeamGetEntity;
	e->Mod_SampleSizeForFace = xh_Mod_SampleSizeForFace;
	e->Mod_BoxVisible = xh_Mod_BoxVisible;
	e->Mod_PointInLeaf = xh_Mod_PointInLeaf;
	e->CL_DrawEFX = xh_CL_DrawEFX;
	e->CL_ThinkParticle = xh_CL_ThinkParticle;
	e->R_StoreEfrags = xh_R_StoreEfrags;
	e->Mod_Extradata = xh_Mod_Extradata;
	e->CL_ExtraUpdate = xh_void;
	e->Host_Error = xh_Host_Error;
	e->COM_SetRandomSeed = xh_SetRandomSeed;
	e->COM_RandomFloat = xh_RandomFloat;
	e->COM_RandomLong = xh_RandomLong;
	e->Sys_CheckParm = xh_Sys_CheckParm;
	e->_Mem_AllocPool = xh_Mem_AllocPool;
	e->_Mem_FreePool = xh_Mem_FreePool;
	e->_Mem_Alloc = xh_Mem_Alloc;
	e->_Mem_Realloc = xh_Mem_Realloc;
	e->_Mem_Free = xh_Mem_Free;
	e->R_Init_Video = xh_R_Init_Video;
	e->R_Free_Video = xh_void;
	e->GL_SetAttribute = xh_GL_SetAttribute;
	e->GL_GetAttribute = xh_GL_GetAttribute;
	e->GL_GetProcAddress = xh_GL_GetProcAddress;
	e->GL_SwapBuffers = xh_void;
	e->SW_CreateBuffer = xh_SW_CreateBuffer;
	e->SW_LockBuffer = xh_SW_LockBuffer;
	e->SW_UnlockBuffer = xh_SW_UnlockBuffer;
	e->R_FatPVS = xh_R_FatPVS;
	e->pfnTime = xh_Time;
	e->Image_SetForceFlags = xh_Image_SetForceFlags;
	e->Image_ClearForceFlags = xh_Image_ClearForceFlags;
	e->Image_CustomPalette = xh_Image_CustomPalette;
	e->Image_Process = xh_Image_Process;
	e->FS_LoadImage = xh_FS_LoadImage;
	e->FS_CopyImage = xh_FS_CopyImage;
	e->FS_FreeImage = xh_FS_FreeImage_any;
	e->FS_SaveImage = xh_FS_SaveImage;
	e->Image_GetPFDesc = xh_Image_GetPFDesc;
	e->Image_SetMDLPointer = xh_Image_SetMDLPointer;
	e->EV_GetPhysent = xh_EV_GetPhysent;
	e->SetLocalLightLevel = xh_SetLocalLightLevel;
	e->CL_GetRemapInfoForEntity = (void *)xh_null_ptr;
	e->CL_EntitySetRemapColors = (void *)xh_null_ptr; /* no colour remapping in the harness: returns false */
	e->pfnStudioEvent = (void *)xh_null_ptr;
	e->pfnPlayerInfo = (void *)xh_null_ptr;
	e->pfnGetPlayerState = (void *)xh_null_ptr;
	e->GetDefaultSprite = (void *)xh_GetDefaultSprite;
	e->EV_VisTraceLine = xh_EV_VisTraceLine;
	e->R_GetSpriteFrame = (void *)xh_R_GetSpriteFrame;
	e->R_StudioGetAnim = (void *)xh_R_StudioGetAnim;
	e->CL_AllocElight = (void *)xh_null_ptr;
	e->EV_TraceSurface = (void *)xh_null_ptr;
	e->GetScreenFade = (void *)xh_null_ptr;
	e->pfnDrawNormalTriangles = xh_void;
	e->pfnDrawTransparentTriangles = xh_void;
	e->drawFuncs = &xh.drawfuncs; /* non-NULL struct of NULL hooks (cl_render.c) */
	xh_fsapi.AllowDirectPaths = xh_AllowDirectPaths;
	e->fsapi = &xh_fsapi;
}

/* ------------------------------------------------------------------ */
/* default cvars the engine creates before loading a renderer            */
/* (engine/client/dll_int/ref_common.c:35-44,755-780; engine/client/gamma.c:29-33; ref_api.h:676-703) */

static void xh_default_cvars( void )
{
	static const char *const kv[][2] = {
		{ "gamma", "2.5" }, { "brightness", "0.0" }, { "lightgamma", "2.5" }, { "texgamma", "2.0" }, { "direct", "0.9" },
		{ "r_showtextures", "0" }, { "r_speeds", "0" }, { "r_fullbright", "0" }, { "r_norefresh", "0" },
		{ "r_lightmap", "0" }, { "r_dynamic", "1" }, { "r_drawentities", "1" }, { "r_decals", "4096" },
		{ "r_showhull", "0" }, { "gl_vsync", "1" }, { "gl_clear", "0" }, { "cl_himodels", "1" },
		{ "cl_lightstyle_lerping", "0" }, { "tracerred", "0.8" }, { "tracergreen", "0.8" }, { "tracerblue", "0.4" },
		{ "traceralpha", "0.5" }, { "r_sprite_lerping", "1" }, { "r_sprite_lighting", "1" }, { "r_drawviewmodel", "1" },
		{ "r_glowshellfreq", "2.2" }, { "host_allow_materials", "0" }, { "r_pvs_radius", "0.1" },
		{ "r_lighting_modulate", "0.6" }, { "r_adjust_fov", "1" },
	};
	for( size_t i = 0; i < sizeof( kv ) / sizeof( kv[0] ); i++ )
		xh_cvar_create( kv[i][0], kv[i][1], 0 );
}

/* ------------------------------------------------------------------ */
/* builtin textures the engine creates right after R_Init (ref_common.c:494-545) */

static uint xh_rgba( uint r, uint g, uint b, uint a ) { return r | g << 8 | b << 16 | a << 24; }

static void xh_builtin_textures( void )
{
	uint chk[256], part[64], flat[16];
	for( int y = 0; y < 16; y++ )
		for( int x = 0; x < 16; x++ )
			chk[y * 16 + x] = (( y < 8 ) != ( x < 8 )) ? xh_rgba( 255, 0, 255, 255 ) : xh_rgba( 0, 0, 0, 255 );
	xh.ref.GL_CreateTexture( REF_DEFAULT_TEXTURE, 16, 16, chk, 0 );

	for( int i = 0; i < 64; i++ ) part[i] = xh_rgba( 255, 255, 255, 0 );
	for( int y = 1; y < 5; y++ )
		for( int x = 1; x < 5; x++ )
			if( !(( x == 1 || x == 4 ) && ( y == 1 || y == 4 )))
				part[y * 8 + x] = xh_rgba( 255, 255, 255, 255 );
	xh.ref.GL_CreateTexture( REF_PARTICLE_TEXTURE, 8, 8, part, TF_CLAMP | TF_HAS_ALPHA );

	for( int i = 0; i < 16; i++ ) flat[i] = 0xFFFFFFFFu;
	xh.ref.GL_CreateTexture( REF_WHITE_TEXTURE, 4, 4, flat, 0 );
	for( int i = 0; i < 16; i++ ) flat[i] = xh_rgba( 127, 127, 127, 255 );
	xh.ref.GL_CreateTexture( REF_GRAY_TEXTURE, 4, 4, flat, 0 );
	for( int i = 0; i < 16; i++ ) flat[i] = xh_rgba( 0, 0, 0, 255 );
	xh.ref.GL_CreateTexture( REF_BLACK_TEXTURE, 4, 4, flat, 0 );
}

/* ------------------------------------------------------------------ */

#include <execinfo.h>
#include <signal.h>
#include <unistd.h>
/* XH_BACKTRACE in the environment: a fault inside a renderer prints its frames (no debugger on the boxes) */
static void xh_on_fault( int sig )
{
	void *frames[48];
	int n = backtrace( frames, 48 );
	dprintf( 2, "[xh_host] signal %d\n", sig );
	backtrace_symbols_fd( frames, n, 2 );
	_exit( 128 + sig );
}

int xh_init( const char *ref_path, int width, int height )
{
	if( getenv( "XH_BACKTRACE" )) { signal( SIGSEGV, xh_on_fault ); signal( SIGBUS, xh_on_fault ); signal( SIGFPE, xh_on_fault ); }
	if( xh.dll ) { snprintf( xh_errbuf, sizeof( xh_errbuf ), "already initialised" ); return -1; }

	xh.verbose = getenv( "XH_VERBOSE" ) != NULL;
	xh.dll = dlopen( ref_path, RTLD_NOW | RTLD_LOCAL );
	if( !xh.dll ) { snprintf( xh_errbuf, sizeof( xh_errbuf ), "dlopen: %s", dlerror()); return -2; }

	REFAPI get = (REFAPI)dlsym( xh.dll, GET_REF_API );
	if( !get ) { snprintf( xh_errbuf, sizeof( xh_errbuf ), "no GetRefAPI in %s", ref_path ); return -3; }

	xh_default_cvars();
	xh_build_gamma( xh_cvar_get( "gamma" ), xh_cvar_get( "brightness" ), xh_cvar_get( "texgamma" ), xh_cvar_get( "lightgamma" ));

	memset( &xh.refState, 0, sizeof( xh.refState ));
	xh.refState.width = width; xh.refState.height = height;
	xh.refState.scale_x = xh.refState.scale_y = 1.0f;
	xh.refState.desktopBitsPixel = 16;
	xh.cl.maxclients = 1; xh.cl.playernum = 0; xh.cl.viewentity = 1;
	xh.cl.time = 1.0; xh.cl.oldtime = 1.0 - 1.0 / 60.0;
	xh.host.realtime = 1.0; xh.host.frametime = 1.0 / 60.0;
	xh.movevars.zmax = 4096.0f;
	for( int i = 0; i < 256; i++ )
		xh.palette[i].r = xh.palette[i].g = xh.palette[i].b = (byte)i;
	xh.max_entities = XH_MAX_ENTITIES;
	xh.entities = calloc( xh.max_entities, sizeof( cl_entity_t ));
	xh.rng = 0xB200;

	ref_api_t eng;
	xh_fill_engfuncs( &eng ); /* the engine passes a private copy (ref_common.c:549,571) */
	if( get( REF_API_VERSION, &xh.ref, &eng, &xh.refState ) != REF_API_VERSION )
	{
		snprintf( xh_errbuf, sizeof( xh_errbuf ), "GetRefAPI version mismatch" );
		return -4;
	}
	if( !xh.ref.R_Init())
	{
		snprintf( xh_errbuf, sizeof( xh_errbuf ), "R_Init failed" );
		return -5;
	}
	xh_builtin_textures();
	xh.ref.R_ProcessEntData( true, xh.entities, xh.max_entities );
	/* CL_InitStudioAPI (engine/client/dll_int/cl_game.c:3966): no client DLL overrides the studio interface here, the
	 * renderer's default one is installed */
	{
		static engine_studio_api_t studio_api;
		static r_studio_interface_t default_draw;
		if( xh.ref.R_StudioFillAPI( &studio_api, &default_draw ))
			xh.ref.R_StudioSetDrawInterface( &default_draw );
	}
	return 0;
}

void xh_shutdown( void )
{
	if( !xh.dll ) return;
	xh.ref.R_Shutdown();
	/* the library is left mapped: renderer globals die with the process */
	xh.dll = NULL;
}

/* R_Shutdown + R_Init on the loaded renderer (what the engine's vid_restart does); the map has to be loaded again */
int xh_restart( void )
{
	if( !xh.dll )
		return -1;
	xh.ref.R_Shutdown();
	return xh.ref.R_Init() ? 0 : -5;
}

void *xh_ref_symbol( const char *name ) { return xh.dll ? dlsym( xh.dll, name ) : NULL; }
uint64_t xh_callback_count( void ) { return xh.callbacks; }

/* ------------------------------------------------------------------ */
/* time / light styles                                                   */

/* the renderer's R_LightPoint slot at p -> r, g, b, a */
void xh_light_point( const float *p, int *out )
{
	colorVec c = xh.ref.R_LightPoint( p );
	out[0] = c.r; out[1] = c.g; out[2] = c.b; out[3] = c.a;
}

/* V_CheckGamma (engine/client/gamma.c:117-141): new tables, then the renderer is told */
void xh_gamma_changed( float gamma, float brightness, float texgamma, float lightgamma )
{
	xh_build_gamma( gamma, brightness, texgamma, lightgamma );
	if( xh.ref.R_GammaChanged )
		xh.ref.R_GammaChanged( false );
}

void xh_set_time( double time, double oldtime )
{
	xh.cl.time = time; xh.cl.oldtime = oldtime;
	xh.host.realtime = time; xh.host.frametime = time - oldtime;
}

/* engine/client/cl_tent.c:2390-2424 CL_SetLightstyle */
void xh_set_lightstyle( int style, const char *pattern, float time )
{
	if( style < 0 || style >= MAX_LIGHTSTYLES ) return;
	lightstyle_t *ls = &xh.lightstyles[style];
	memset( ls, 0, sizeof( *ls ));
	strncpy( ls->pattern, pattern, sizeof( ls->pattern ) - 1 );
	ls->length = (int)strlen( ls->pattern );
	ls->time = time;
	for( int i = 0; i < ls->length; i++ ) ls->map[i] = (float)( ls->pattern[i] - 'a' );
	ls->interp = ls->length > 1;
	for( int i = 0; i < ls->length - 1; i++ )
		if( fabsf( ls->map[i] - ls->map[( i + 1 ) % ls->length] ) > 3.0f ) { ls->interp = false; break; }
}

void xh_run_lightstyles( void ) { xh.ref.CL_RunLightStyles( xh.lightstyles ); }

/* ------------------------------------------------------------------ */
/* frames                                                                */

void xh_begin_frame( void ) { xh.ref.R_BeginFrame( false ); }
void xh_end_frame( void ) { xh.ref.R_EndFrame(); }

void xh_render_view( const float *origin, const float *angles, float fov_x, float fov_y, int x, int y, int w, int h, int flags )
{
	ref_viewpass_t rvp;
	memset( &rvp, 0, sizeof( rvp ));
	rvp.viewport[0] = x; rvp.viewport[1] = y; rvp.viewport[2] = w; rvp.viewport[3] = h;
	for( int i = 0; i < 3; i++ ) { rvp.vieworigin[i] = origin[i]; rvp.viewangles[i] = angles[i]; }
	rvp.viewentity = xh.cl.viewentity;
	rvp.fov_x = fov_x; rvp.fov_y = fov_y;
	rvp.flags = flags;
	for( int i = 0; i < 3; i++ ) { xh.refState.vieworg[i] = origin[i]; xh.refState.viewangles[i] = angles[i]; }
	xh.ref.GL_RenderFrame( &rvp );
}

/* the engine's frame loop for one view per frame: R_BeginFrame, GL_RenderFrame, R_EndFrame; the software surface
 * (display format) is copied out after each frame.  Works with any renderer plugin. */
int xh_frame_views( const float *cams, int n, float fov_x, float fov_y, int x, int y, int w, int h, int flags, void *display_out )
{
	for( int i = 0; i < n; i++ )
	{
		const float *c = cams + 6 * i;
		xh.ref.R_BeginFrame( false );
		xh_render_view( c, c + 3, fov_x, fov_y, x, y, w, h, flags );
		xh.ref.R_EndFrame();
		if( !xh.swbuffer )
			return -1;
		size_t bytes = (size_t)xh.sw_w * xh.sw_h * xh.sw_bpp;
		memcpy((char *)display_out + bytes * i, xh.swbuffer, bytes );
	}
	return 0;
}

/* a HUD picture: an RGBA image handed to the renderer (the engine's pfnSPR_Load / Draw_* paths end in GL_LoadTexture*) */
int xh_load_pic( const char *name, int w, int h, const void *rgba, int texflags )
{
	return xh.ref.GL_CreateTexture( name, w, h, rgba, texflags );
}

/* one frame as the engine drives it, with a HUD pass: R_BeginFrame, R_Set2DMode( false ) + the view (V_RenderView),
 * R_Set2DMode( true ) + the 2-D draws (V_PostRender / SCR_UpdateScreen), R_EndFrame.
 * hud: nhud x 15 floats {kind 0 pic / 1 fill, rendermode, r, g, b, a, x, y, w, h, s1, t1, s2, t2, texnum} */
int xh_frame_views_hud( const float *cams, int n, float fov_x, float fov_y, int x, int y, int w, int h, int flags,
	const float *hud, int nhud, void *display_out )
{
	for( int i = 0; i < n; i++ )
	{
		const float *c = cams + 6 * i;
		xh.ref.R_BeginFrame( false );
		xh.ref.R_Set2DMode( false );
		xh_render_view( c, c + 3, fov_x, fov_y, x, y, w, h, flags );
		xh.ref.R_Set2DMode( true );
		for( int k = 0; k < nhud; k++ )
		{
			const float *q = hud + 15 * k;
			if( (int)q[0] == 1 )
				xh.ref.FillRGBA( (int)q[1], q[6], q[7], q[8], q[9], (byte)q[2], (byte)q[3], (byte)q[4], (byte)q[5] );
			else
			{
				xh.ref.GL_SetRenderMode( (int)q[1] );
				xh.ref.Color4ub( (byte)q[2], (byte)q[3], (byte)q[4], (byte)q[5] );
				xh.ref.R_DrawStretchPic( q[6], q[7], q[8], q[9], q[10], q[11], q[12], q[13], (int)q[14] );
			}
		}
		xh.ref.R_EndFrame();
		if( !xh.swbuffer )
			return -1;
		size_t bytes = (size_t)xh.sw_w * xh.sw_h * xh.sw_bpp;
		memcpy((char *)display_out + bytes * i, xh.swbuffer, bytes );
	}
	return 0;
}

/* an entry of the engine's dynamic light table (CL_AllocDlight, engine/client/cl_efx.c); radius 0 switches it off */
int xh_set_dlight( int i, const float *origin, float radius, int r, int g, int b, float die, float minlight )
{
	if( i < 0 || i >= MAX_DLIGHTS )
		return -1;
	dlight_t *l = &xh.dlights[i];
	memset( l, 0, sizeof( *l ));
	for( int k = 0; k < 3; k++ ) l->origin[k] = origin[k];
	l->radius = radius; l->die = die; l->minlight = minlight;
	l->color.r = (uint8_t)r; l->color.g = (uint8_t)g; l->color.b = (uint8_t)b;
	l->key = i + 1;
	return 0;
}

void xh_clear_scene( void ) { xh.ref.R_ClearScene(); xh.num_entities = 1; }

/* ------------------------------------------------------------------ */
/* sprite and studio models (engine/common/mod_sprite.c, mod_studio.c)  */

static mspriteframe_t *xh_R_GetSpriteFrame( const model_t *pModel, int frame, float yaw )
{
	(void)yaw;
	msprite_t *psprite = pModel ? (msprite_t *)pModel->cache.data : NULL;
	if( !psprite ) return NULL;
	if( frame < 0 || frame >= psprite->numframes ) frame = 0;
	return psprite->frames[frame].frameptr; /* SPR_SINGLE frames only */
}

static model_t *xh_new_model( const char *name, modtype_t type )
{
	if( xh.cl.nummodels + 1 >= MAX_MODELS ) return NULL;
	model_t *m = calloc( 1, sizeof( *m ));
	strncpy( m->name, name, sizeof( m->name ) - 1 );
	m->type = type;
	m->needload = 1; /* NL_PRESENT */
	xh.cl.models[++xh.cl.nummodels] = m;
	return m;
}

/* engine/common/mod_studio.c R_StudioGetAnim: sequence group 0 lives in the model file itself (demand-loaded groups
 * are not part of this harness) */
static void *xh_R_StudioGetAnim( studiohdr_t *hdr, model_t *mod, mstudioseqdesc_t *pseqdesc )
{
	(void)mod;
	if( !hdr || !pseqdesc || pseqdesc->seqgroup != 0 ) return NULL;
	return (byte *)hdr + pseqdesc->animindex;
}

/* a studio model from an MDL v10 image in memory: what Mod_LoadStudioModel (engine/common/mod_studio.c:1126) leaves in
 * model_t - a private copy of the file as cache.data, textures handed to the renderer, bounds from the header */
int xh_load_studio( const char *name, const void *buffer, size_t len )
{
	const studiohdr_t *in = (const studiohdr_t *)buffer;
	if( !buffer || len < sizeof( studiohdr_t ) || in->ident != IDSTUDIOHEADER || in->version != STUDIO_VERSION || (size_t)in->length > len )
		return -1;
	model_t *m = xh_new_model( name, mod_studio );
	if( !m ) return -1;
	m->cache.data = malloc( in->length );
	memcpy( m->cache.data, buffer, in->length );
	studiohdr_t *phdr = (studiohdr_t *)m->cache.data;
	xh.ref.Mod_StudioLoadTextures( m, phdr );
	for( int i = 0; i < 3; i++ ) { m->mins[i] = phdr->bbmin[i]; m->maxs[i] = phdr->bbmax[i]; }
	m->numframes = 1;
	m->flags = phdr->flags;
	xh.ref.Mod_ProcessRenderData( m, true, buffer, len );
	return xh.cl.nummodels;
}

/* a one-frame sprite model from an RGBA image: what Mod_LoadSpriteModel leaves in model_t (msprite_t with SPR_SINGLE
 * frames whose textures went through the renderer's GL_LoadTexture); returns the model index */
int xh_create_sprite( const char *name, int w, int h, const void *rgba, int type, int texformat, int texflags )
{
	model_t *m = xh_new_model( name, mod_sprite );
	if( !m ) return -1;
	msprite_t *ps = calloc( 1, sizeof( *ps ));
	mspriteframe_t *fr = calloc( 1, sizeof( *fr ));
	ps->type = (short)type; ps->texFormat = (short)texformat;
	ps->maxwidth = w; ps->maxheight = h; ps->numframes = 1;
	ps->radius = (int)sqrt((double)(( w >> 1 ) * ( w >> 1 ) + ( h >> 1 ) * ( h >> 1 )));
	fr->width = w; fr->height = h;
	fr->up = (float)( h / 2 ); fr->left = (float)( -( w / 2 )); fr->down = fr->up - h; fr->right = w + fr->left;
	char texname[128];
	snprintf( texname, sizeof( texname ), "#%s(frame0)", name );
	fr->gl_texturenum = xh.ref.GL_CreateTexture( texname, w, h, rgba, texflags );
	ps->frames[0].type = SPR_SINGLE; ps->frames[0].frameptr = fr;
	m->cache.data = ps;
	m->mins[0] = m->mins[1] = -w * 0.5f; m->maxs[0] = m->maxs[1] = w * 0.5f;
	m->mins[2] = -h * 0.5f; m->maxs[2] = h * 0.5f;
	m->numframes = 1;
	xh.ref.Mod_ProcessRenderData( m, true, NULL, 0 );
	return xh.cl.nummodels;
}

/* ip: modelindex, rendermode, renderamt, r, g, b, renderfx, frame, sequence, body, skin, movetype, aiment; fp: origin[3], angles[3], scale, framerate, animtime */
int xh_add_entity( const int *ip, const float *fp )
{
	const int modelindex = ip[0];
	if( modelindex <= 0 || modelindex > xh.cl.nummodels || !xh.cl.models[modelindex] ) return -1;
	if( xh.num_entities >= (int)xh.max_entities ) return -1;
	if( xh.num_entities < 1 ) xh.num_entities = 1; /* entity 0 is the world */
	int idx = xh.num_entities++;
	cl_entity_t *e = &xh.entities[idx];
	memset( e, 0, sizeof( *e ));
	e->index = idx;
	e->model = xh.cl.models[modelindex];
	e->curstate.modelindex = modelindex;
	e->curstate.rendermode = ip[1]; e->curstate.renderamt = ip[2];
	e->curstate.rendercolor.r = (byte)ip[3]; e->curstate.rendercolor.g = (byte)ip[4]; e->curstate.rendercolor.b = (byte)ip[5];
	e->curstate.renderfx = ip[6]; e->curstate.frame = (float)ip[7];
	e->curstate.sequence = ip[8]; e->curstate.body = ip[9]; e->curstate.skin = ip[10];
	e->curstate.movetype = ip[11]; e->curstate.aiment = ip[12];
	e->curstate.scale = fp[6]; e->curstate.framerate = fp[7]; e->curstate.animtime = fp[8];
	for( int i = 0; i < 3; i++ )
	{
		e->origin[i] = e->curstate.origin[i] = e->baseline.origin[i] = fp[i];
		e->angles[i] = e->curstate.angles[i] = fp[3 + i];
		e->latched.prevorigin[i] = fp[i]; e->latched.prevangles[i] = fp[3 + i];
	}
	e->latched.prevanimtime = e->curstate.animtime;
	e->latched.prevsequence = e->curstate.sequence;
	return xh.ref.R_AddEntity( e, ET_NORMAL ) ? idx : -2;
}

int xh_add_brush_entity( int modelindex, const float *origin, const float *angles, int rendermode, int renderamt )
{
	if( modelindex <= 1 || modelindex > xh.cl.nummodels || !xh.cl.models[modelindex] ) return -1;
	if( xh.num_entities >= (int)xh.max_entities ) return -1;
	int idx = xh.num_entities++;
	cl_entity_t *e = &xh.entities[idx];
	memset( e, 0, sizeof( *e ));
	e->index = idx;
	e->model = xh.cl.models[modelindex];
	e->curstate.modelindex = modelindex;
	e->curstate.rendermode = rendermode;
	e->curstate.renderamt = renderamt;
	for( int i = 0; i < 3; i++ )
	{
		e->origin[i] = e->curstate.origin[i] = origin[i];
		e->angles[i] = e->curstate.angles[i] = angles[i];
	}
	return xh.ref.R_AddEntity( e, ET_NORMAL ) ? idx : -1;
}
