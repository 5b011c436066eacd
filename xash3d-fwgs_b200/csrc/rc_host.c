/*
 * rc_host.c - ref/cuda: a renderer plugin for Xash3D FWGS that exports the GetRefAPI /
 * ref_interface_t surface of engine/ref_api.h (:527-661) and renders batched camera views of BSP30
 * worlds with sm_100a CUDA kernels (rc_device.cu).  Drop-in for ref/soft (`-ref cuda`).
 *
 * This file is the C host side: the 78-slot function table, the texture table (pixel_t encoding of
 * ref/soft/r_image.c:214-277 restated), the blend LUT builder (r_glblit.c:411-470), the flattening of
 * the engine's model_t into device arrays at R_NewMap, the per-view constants of
 * R_SetupFrustum/R_SetupFrameQ/R_ViewChanged (r_main.c:387, r_misc.c:101-322), light styles
 * (ref/common/ref_light.c:30-87) and the view queue.  All pixels are produced on the GPU; there is no
 * CPU rasteriser in this library.
 */
#define REF_DLL 1
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <strings.h>

#include "build.h"
#include "port.h"
#include "xash3d_types.h"
#include "const.h"
#include "cvardef.h"
#include "com_model.h"
#include "cl_entity.h"
#include "ref_api.h"
#include "ref_params.h"
#include "enginefeatures.h"
#include "com_strings.h"
#include "lightstyle.h"
#include "xash3d_mathlib.h"
#include "sprite.h"
#include "studio.h"
#include "triangleapi.h"
#include "entity_types.h"
#include "pmove.h"
#include "pm_defs.h"
#include "protocol.h"

#include "rc_dev.h"
#include "ref_cuda.h"

#define RC_MAX_TEXTURES 8192
#define RC_MAX_QUEUED_VIEWS 16384
#define RC_MAX_VISIBLE 2048 /* MAX_VISIBLE_PACKET */
#define LM_SAMPLE_SIZE 16

DEFINE_ENGINE_SHARED_CVAR_LIST()

ref_api_t      gEngfuncs;
ref_globals_t *gpGlobals;
static ref_client_t *gp_cl;
static ref_host_t   *gp_host;

static CVAR_DEFINE_AUTO( sw_clearcolor, "48999", 0, "screen clear color" );
static CVAR_DEFINE_AUTO( sw_mipcap, "0", FCVAR_GLCONFIG, "" );
static CVAR_DEFINE_AUTO( sw_texfilt, "0", FCVAR_GLCONFIG, "texture dither" );
static CVAR_DEFINE_AUTO( sw_mipscale, "1", FCVAR_GLCONFIG, "nothing" );
static CVAR_DEFINE_AUTO( sw_noalphabrushes, "0", FCVAR_GLCONFIG, "" );
static CVAR_DEFINE_AUTO( cu_device, "0", FCVAR_GLCONFIG, "CUDA device ordinal used by ref_cuda" );
static CVAR_DEFINE_AUTO( r_novis, "0", 0, "" );
static CVAR_DEFINE_AUTO( r_traceglow, "0", FCVAR_GLCONFIG, "cull flares behind models" ); /* ref/soft/r_main.c:69 */
/* ref/common/ref_light.c:21 */
static CVAR_DEFINE_AUTO( r_dlight_virtual_radius, "3", FCVAR_GLCONFIG, "increase dlight radius virtually by this amount" );

/* ---------------------------------------------------------------------------------------------- */
/* texture table (image_t of ref/soft/r_local.h:287-312, host mirror; texels are uploaded as one arena) */

typedef struct
{
	char      name[256];
	int       width, height, srcWidth, srcHeight;
	uint      flags;
	uint32_t *rgba;      /* the RGBA8 source as it was handed over: the device encodes pixel_t mips and the alpha plane from it */
	qboolean  has_alpha; /* an alpha plane exists (TF_HAS_ALPHA): alpha_pixels of the reference */
	qboolean  colorkey;  /* IMAGE_ONEBIT_ALPHA source and sw_noalphabrushes off when it was loaded */
} rc_image_t;
#define RC_IMAGE_LIVE( t ) (( t )->rgba != NULL )

static struct
{
	qboolean     initialized;
	poolhandle_t pool;
	rc_image_t   images[RC_MAX_TEXTURES];
	int          numImages;
	qboolean     textures_dirty;

	uint16_t     colormap[32 * 8192];
	int32_t      sintable[1280];
	const uint16_t *lightgamma; /* engine table */
	uint16_t lightgamma_sent[1024]; /* what the device has */

	int          lightstylevalue[MAX_LIGHTSTYLES];

	/* world */
	model_t     *world;
	qboolean     scene_valid;
	int32_t     *surf_texnum;
	int          maxdepth;

	/* view queue between R_BeginFrame and R_EndFrame */
	rc_view_t   *views;
	int          nviews, maxviews;
	struct rc_glowsrc_s { vec3_t origin, color; float scale, blend; int renderfx; } *glowsrc; /* the batch's kRenderGlow sprites */
	int          nglow, maxglow, ntracers;
	rc_glow_t    *glowrecs; int maxglowrecs;
	ref_viewpass_t last_rvp;      /* the view set up last: the WorldToScreen slot answers for it */
	qboolean     batch_wvp;       /* the batch draws tracers or glow sprites: its views carry RI.worldviewProjectionMatrix */
	int          particleTexture; /* tr.particleTexture: the engine's "*particle" builtin, remembered by name (r_image.c:646) */
	int          screen_view, screen_valid; /* the frame of the last batch R_EndFrame showed (VID_ScreenShot) */
	rc_cmd2d_t  *cmds2d; int ncmds2d, maxcmds2d; /* the frame's HUD pass (R_DrawStretchPic / FillRGBA), applied in R_EndFrame */

	void        *swbuffer_info_valid;
	int          had_dlights;
	uint         sw_stride, sw_bpp, sw_r, sw_g, sw_b;
	uint         map_valid, map_bpp, map_r, map_g, map_b; /* display format the device's screen map was built for */
	unsigned long long last_h2d;
	char         error[512];

	/* entity lists of the frame (draw_list_t, ref/soft/r_local.h; R_AddEntity, r_main.c:311) */
	cl_entity_t *solid_ents[RC_MAX_VISIBLE], *trans_ents[RC_MAX_VISIBLE], *edge_ents[RC_MAX_VISIBLE];
	int          num_solid, num_trans, num_edge;
	cl_entity_t *entities; uint max_entities;

	/* the immediate-mode state TriAPI / GL_Bind keep in viddef_t and r_triapi.c's statics */
	struct
	{
		int      rendermode;      /* vid.rendermode */
		int      is2d;
		uint     alpha, color;    /* vid.alpha, vid.color */
		uint     light;
		short    s, t;
		int      prim, vertcount, n;
		int      drawer, texnum, tex_ofs, skinwidth, skinheight; /* r_affinetridesc + d_pdrawspans as GL_Bind left them */
		float    triRGBA[4];
		int      triRenderMode;
		rc_corner_t triv[3];
		float    blend;           /* tr.blend */
	} tri;

	/* draw items of the batch being prepared */
	rc_texture_t *textab; int ntextab; /* the texture table as last sent to the device (arena offsets) */
	rc_item_t   *items; int nitems, maxitems;
	rc_corner_t *corners; int ncorners, maxcorners;
	rc_sortent_t *sortents; int nsortents, maxsortents;
	int          cur_sortent;
	qboolean     warned[8];
} rc;

/* studio models of the batch being prepared + the static per-submodel tables (see the studio section) */
static struct
{
	byte   *arena; size_t used, cap; qboolean dirty;
	struct { const void *hdr, *sub; int skin; int32_t sverts, svertbone, snorms, snormbone, snormflags; int nmesh; int32_t mesh_first[MAXSTUDIOMESHES], mesh_ntris[MAXSTUDIOMESHES], mesh_flags[MAXSTUDIOMESHES], mesh_tex[MAXSTUDIOMESHES]; } sub[256];
	int     nsub;
	rc_spart_t *parts; int nparts, maxparts;
	float  *bones; int nbones, maxbones;      /* matrix3x4 per bone */
	float  *blight; int maxblight;            /* vec3 per bone */
	int     totalverts, totalnorms;
} rs;

static void rc_strncpy( char *dst, const char *src, size_t n ) { strncpy( dst, src, n - 1 ); dst[n - 1] = 0; }

static void rc_error( const char *fmt, ... )
{
	va_list ap;
	va_start( ap, fmt );
	vsnprintf( rc.error, sizeof( rc.error ), fmt, ap );
	va_end( ap );
	gEngfuncs.Con_Printf( S_ERROR "ref_cuda: %s\n", rc.error );
}

#define RC_CHECK_DEV( call ) do { if(( call ) != 0 ) gEngfuncs.Host_Error( "ref_cuda: %s failed: %s\n", #call, rc_dev_last_error()); } while( 0 )

static void *rc_malloc( size_t n ) { return gEngfuncs._Mem_Alloc( rc.pool, n, true, __FILE__, __LINE__ ); }
static void rc_free( void *p ) { if( p ) gEngfuncs._Mem_Free( p, __FILE__, __LINE__ ); }

/* pixel_t code of an RGB triple already reduced to 5/6/5 bits: major RGB332 << 8 | minor GBRGBRGB
 * (ref/soft/r_image.c:261-271, r_glblit.c:463-468) */
static uint16_t rc_encode565( uint r, uint g, uint b )
{
	uint major = ((( r >> 2 ) & 7 ) << 5 ) | ((( g >> 3 ) & 7 ) << 2 ) | (( b >> 3 ) & 3 );
	uint minor = ((( r >> 1 ) & 1 ) << 5 ) | (( r & 1 ) << 2 )
		| ((( g >> 2 ) & 1 ) << 7 ) | ((( g >> 1 ) & 1 ) << 4 ) | (( g & 1 ) << 1 )
		| ((( b >> 2 ) & 1 ) << 6 ) | ((( b >> 1 ) & 1 ) << 3 ) | ( b & 1 );
	return (uint16_t)( major << 8 | ( minor & 0xFF ));
}

/* the 565 triple a pixel_t code stands for (inverse of the above, r_glblit.c:447-452) */
static void rc_decode565( uint color, uint *r, uint *g, uint *b )
{
	uint m = color >> 8, j = color & 0xFF;
	*r = ((( m >> 5 ) << 2 ) & 31 ) | ((( j >> 5 ) & 1 ) << 1 ) | (( j >> 2 ) & 1 );
	*g = ((( m >> 2 ) << 3 ) & 63 ) | ((( j >> 7 ) & 1 ) << 2 ) | ((( j >> 4 ) & 1 ) << 1 ) | (( j >> 1 ) & 1 );
	*b = (( m << 3 ) & 31 ) | ((( j >> 6 ) & 1 ) << 2 ) | ((( j >> 3 ) & 1 ) << 1 ) | ( j & 1 );
}

/* vid.colormap of R_BuildBlendMaps (r_glblit.c:440-470): 32 light levels x 8192 colour codes (pixel >> 3) */
static void rc_build_colormap( void )
{
	for( uint i = 0; i < 8192; i++ )
	{
		uint r1, g1, b1;
		rc_decode565( i << 3, &r1, &g1, &b1 );
		for( uint j = 0; j < 32; j++ )
			rc.colormap[( j << 13 ) | i] = rc_encode565( r1 * j / 32, g1 * j / 32, b1 * j / 32 );
	}
}

static rc_image_t *rc_find_image( const char *name )
{
	for( int i = 0; i < rc.numImages; i++ )
		if( rc.images[i].name[0] && !strcasecmp( rc.images[i].name, name ))
			return &rc.images[i];
	return NULL;
}

static rc_image_t *rc_alloc_image( const char *name, uint flags )
{
	int i;
	for( i = 0; i < rc.numImages; i++ )
		if( !rc.images[i].name[0] )
			break;
	if( i == rc.numImages )
	{
		if( rc.numImages == RC_MAX_TEXTURES )
			gEngfuncs.Host_Error( "%s: MAX_TEXTURES limit exceeds\n", __func__ );
		rc.numImages++;
	}
	rc_image_t *tex = &rc.images[i];
	memset( tex, 0, sizeof( *tex ));
	rc_strncpy( tex->name, name, sizeof( tex->name ));
	tex->flags = flags;
	return tex;
}

static void rc_delete_image( rc_image_t *tex )
{
	rc_free( tex->rgba );
	memset( tex, 0, sizeof( *tex ));
	rc.textures_dirty = true;
}

/*
 * Texture intake.  What GL_UploadTexture / GL_BuildMipMap do on the CPU in the reference (ref/soft/r_image.c:125-283: four
 * box-filtered mips encoded to pixel_t, the alpha plane, the colour key) is a kernel here (k_tex_encode, rc_core.h A19): the
 * host only keeps the RGBA8 source of every texture and, when the set changed, sends all sources in one arena
 * (rc_sync_textures).  One routine serves the three loader slots of the interface.
 */
static qboolean rc_name_ok( const char *name )
{
	return name && name[0] && strlen( name ) < sizeof( rc.images->name );
}

static int rc_image_index( const rc_image_t *tex ) { return tex ? (int)( tex - rc.images ) : 0; }

/* tex takes a private copy of pic's pixels; false = format not supported */
static qboolean rc_image_take( rc_image_t *tex, const rgbdata_t *pic )
{
	tex->srcWidth = tex->width = Q_max( 1, pic->width );
	tex->srcHeight = tex->height = Q_max( 1, pic->height );
	if( FBitSet( pic->flags, IMAGE_HAS_ALPHA ))
		SetBits( tex->flags, TF_HAS_ALPHA );
	if( !pic->buffer )
		return true; /* a sized, empty texture */
	if( pic->type != PF_RGBA_32 )
	{
		rc_error( "texture %s: unsupported pixel format %d", tex->name, pic->type );
		return false;
	}
	const size_t bytes = (size_t)tex->width * tex->height * 4;
	rc_free( tex->rgba );
	tex->rgba = rc_malloc( bytes );
	memcpy( tex->rgba, pic->buffer, bytes );
	tex->has_alpha = FBitSet( tex->flags, TF_HAS_ALPHA ) != 0;
	tex->colorkey = tex->has_alpha && !sw_noalphabrushes.value && FBitSet( pic->flags, IMAGE_ONEBIT_ALPHA );
	rc.textures_dirty = true;
	return true;
}

/* register (or, with update, refresh) the texture `name` from pic; 0 = failure, else its number */
static int rc_texture_store( const char *name, const rgbdata_t *pic, uint flags, qboolean update )
{
	rc_image_t *tex = rc_find_image( name );
	if( tex && !update )
		return rc_image_index( tex ); /* known already */
	if( !pic )
		return 0;
	if( update && !tex )
		gEngfuncs.Host_Error( "%s: couldn't find texture %s for update\n", __func__, name );
	if( tex ) SetBits( tex->flags, flags );
	else tex = rc_alloc_image( name, flags );
	if( rc_image_take( tex, pic ))
		return rc_image_index( tex );
	rc_free( tex->rgba );
	memset( tex, 0, sizeof( *tex ));
	return 0;
}

static int GL_LoadTextureFromBuffer( const char *name, rgbdata_t *pic, texFlags_t flags, qboolean update )
{
	return rc_name_ok( name ) ? rc_texture_store( name, pic, flags, update ) : 0;
}

static int GL_LoadTexture( const char *name, const byte *buf, size_t size, int flags )
{
	if( !rc_name_ok( name ))
		return 0;
	const int known = rc_image_index( rc_find_image( name ));
	if( known )
		return known;
	/* the engine's image loader decodes the file; these two switches are what the reference asks of it (r_image.c:508-514) */
	const uint force = ( FBitSet( flags, TF_NOFLIP_TGA ) ? IL_DONTFLIP_TGA : 0 )
		| (( FBitSet( flags, TF_KEEP_SOURCE ) && !FBitSet( flags, TF_EXPAND_SOURCE )) ? IL_KEEP_8BIT : 0 );
	gEngfuncs.Image_SetForceFlags( force );
	rgbdata_t *pic = gEngfuncs.FS_LoadImage( name, buf, size );
	if( !pic )
		return 0;
	const int texnum = rc_texture_store( name, pic, flags, false );
	gEngfuncs.FS_FreeImage( pic );
	return texnum;
}

static int GL_CreateTexture( const char *name, int width, int height, const void *buffer, texFlags_t flags )
{
	if( FBitSet( flags, TF_CUBEMAP ))
		return 0;
	rgbdata_t pic;
	memset( &pic, 0, sizeof( pic ));
	pic.width = width; pic.height = height;
	pic.type = PF_RGBA_32;
	pic.size = width * height * 4;
	pic.flags = IMAGE_HAS_COLOR | ( FBitSet( flags, TF_HAS_ALPHA ) ? IMAGE_HAS_ALPHA : 0 );
	pic.buffer = (byte *)buffer; /* only read: the texture keeps a copy */
	const int texnum = GL_LoadTextureFromBuffer( name, &pic, flags, false );
	if( name && !strcmp( name, REF_PARTICLE_TEXTURE ))
		rc.particleTexture = texnum;
	return texnum;
}

static int GL_FindTexture( const char *name )
{
	rc_image_t *tex = name ? rc_find_image( name ) : NULL;
	return tex ? tex - rc.images : 0;
}

static const char *GL_TextureName( unsigned int texnum )
{
	return texnum < RC_MAX_TEXTURES ? rc.images[texnum].name : "";
}

static const byte *GL_TextureData( unsigned int texnum ) { (void)texnum; return NULL; }

static void GL_FreeTexture( unsigned int texnum )
{
	if( texnum == 0 || texnum >= RC_MAX_TEXTURES || !rc.images[texnum].name[0] )
		return;
	rc_delete_image( &rc.images[texnum] );
}

/* lay out the pixel_t arena (four mips + alpha plane per live texture), gather the RGBA sources in one upload arena and let
 * the device encode them */
static void rc_decals_texels_moved( void );
static void rc_sync_textures( void )
{
	if( !rc.textures_dirty )
		return;
	rc_decals_texels_moved(); /* decal records hold texel arena offsets */
	const int n = rc.numImages;
	rc_free( rc.textab );
	rc_texture_t *tab = rc.textab = rc_malloc( sizeof( rc_texture_t ) * Q_max( 1, n ));
	rc_texsrc_t *src = rc_malloc( sizeof( rc_texsrc_t ) * Q_max( 1, n ));
	rc.ntextab = n;
	size_t ntexels = 0, nrgba = 0;
	for( int i = 0; i < n; i++ )
	{
		const rc_image_t *t = &rc.images[i];
		rc_texture_t *d = &tab[i];
		d->width = t->width; d->height = t->height; d->flags = t->flags; d->alpha_ofs = -1;
		for( int j = 0; j < 4; j++ ) d->ofs[j] = RC_IMAGE_LIVE( t ) ? 0 : -1;
		if( !RC_IMAGE_LIVE( t ))
			continue;
		for( int j = 0; j < 4; j++ )
		{
			d->ofs[j] = (int32_t)ntexels;
			ntexels += (size_t)Q_max( 1, t->width >> j ) * Q_max( 1, t->height >> j );
		}
		if( t->has_alpha )
		{
			d->alpha_ofs = (int32_t)ntexels;
			ntexels += (size_t)t->width * t->height;
		}
		src[i].rgba_ofs = (uint32_t)nrgba;
		src[i].flags = t->colorkey ? RC_TEXSRC_COLORKEY : 0;
		nrgba += (size_t)t->width * t->height;
	}
	uint32_t *arena = rc_malloc(( nrgba + 1 ) * sizeof( uint32_t ));
	for( int i = 0; i < n; i++ )
		if( RC_IMAGE_LIVE( &rc.images[i] ))
			memcpy( arena + src[i].rgba_ofs, rc.images[i].rgba, (size_t)rc.images[i].width * rc.images[i].height * 4 );
	RC_CHECK_DEV( rc_dev_upload_textures( tab, src, n, arena, nrgba, ntexels + 64 ));
	rc_free( arena ); rc_free( src );
	/* textures without an image keep offset 0 in the table the other host code reads (rc_texture_entry checks liveness) */
	for( int i = 0; i < n; i++ )
		if( !RC_IMAGE_LIVE( &rc.images[i] ))
			for( int j = 0; j < 4; j++ ) tab[i].ofs[j] = 0;
	rc.textures_dirty = false;
}

/* ---------------------------------------------------------------------------------------------- */
/* world flattening: engine model_t (common/com_model.h) -> rc_scene_t arrays                       */

static int rc_node_ref( const model_t *mod, const mnode_t *n )
{
	if( n->contents < 0 )
		return -1 - (int)((const mleaf_t *)n - mod->leafs );
	return (int)( n - mod->nodes );
}

static int rc_tree_depth( const model_t *mod, const mnode_t *n, int d )
{
	if( n->contents < 0 )
		return d;
	int a = rc_tree_depth( mod, node_child( n, 0, mod ), d + 1 );
	int b = rc_tree_depth( mod, node_child( n, 1, mod ), d + 1 );
	return a > b ? a : b;
}

/* the random numbers "-" (random-tiled) textures choose their tile with, drawn at R_Init (GL_InitRandomTable,
 * ref/common/ref_context.c:46) */
static uint16_t rc_tile_random[MOD_FRAMES][MOD_FRAMES];

/* which texture a world face shows at the client's time (what R_TextureAnimation, ref/soft/r_surf.c:239, answers for the
 * world entity): the frame of the "+0".."+9" chain whose [anim_min, anim_max) window holds the wanted tick; a "-" texture
 * wants the tile its position in texture space draws from the random table */
static int rc_surface_texture( const msurface_t *s )
{
	texture_t *first = s->texinfo->texture;
	if( !first->anim_total )
		return first->gl_texturenum;
	int want;
	if( first->name[0] == '-' )
	{
		const int col = (int)(( s->texturemins[0] + ( first->width << 16 )) / first->width ) % MOD_FRAMES;
		const int row = (int)(( s->texturemins[1] + ( first->height << 16 )) / first->height ) % MOD_FRAMES;
		want = rc_tile_random[col][row] % first->anim_total;
	}
	else
	{
		const int ticks_per_second = FBitSet( rc.images[first->gl_texturenum].flags, TF_QUAKEPAL ) ? 10 : 20;
		want = (int)( gp_cl->time * ticks_per_second ) % first->anim_total;
	}
	int hops = 0;
	for( texture_t *t = first; t && hops <= MOD_FRAMES; t = t->anim_next, hops++ )
		if( want >= t->anim_min && want < t->anim_max )
			return t->gl_texturenum;
	return first->gl_texturenum; /* broken chain */
}

static void rc_resolve_texnums( qboolean force )
{
	model_t *w = rc.world;
	qboolean changed = force;
	for( int i = 0; i < w->numsurfaces; i++ )
	{
		int t = rc_surface_texture( &w->surfaces[i] );
		if( t != rc.surf_texnum[i] ) { rc.surf_texnum[i] = t; changed = true; }
	}
	if( changed && !force )
		RC_CHECK_DEV( rc_dev_update_surface_texnums( rc.surf_texnum, w->numsurfaces ));
}

static void rc_clear_decals( void );
static void R_NewMap( void )
{
	model_t *w = gp_cl->models[1];
	rc.world = w;
	rc_clear_decals(); /* R_ClearDecals, r_main.c:1234 */
	rc.scene_valid = false;
	if( !w )
		return;
	if( FBitSet( w->flags, MODEL_QBSP2 ))
	{
		gEngfuncs.Host_Error( "Sorry, ref_cuda can't load maps in BSP2 format.\n" );
		return;
	}

	int numleafs = w->numleafs + 1; /* model_t.numleafs excludes the shared solid leaf 0 */
	rc_scene_t sc;
	memset( &sc, 0, sizeof( sc ));
	sc.numplanes = w->numplanes; sc.numnodes = w->numnodes; sc.numleafs = numleafs;
	sc.numsurfaces = w->numsurfaces; sc.nummarks = w->nummarksurfaces; sc.numsurfedges = w->numsurfedges;
	sc.numedges = w->numedges; sc.numverts = w->numvertexes;
	sc.numclusters = w->numleafs;
	sc.visbytes = ( w->numleafs + 7 ) >> 3;
	sc.headnode = 0;

	rc_plane_t *planes = rc_malloc( sizeof( rc_plane_t ) * Q_max( 1, w->numplanes ));
	for( int i = 0; i < w->numplanes; i++ )
	{
		VectorCopy( w->planes[i].normal, planes[i].normal );
		planes[i].dist = w->planes[i].dist;
		planes[i].type = w->planes[i].type;
	}

	rc_node_t *nodes = rc_malloc( sizeof( rc_node_t ) * Q_max( 1, w->numnodes ));
	for( int i = 0; i < w->numnodes; i++ )
	{
		const mnode_t *n = &w->nodes[i];
		nodes[i].plane = (int)( n->plane - w->planes );
		nodes[i].child[0] = rc_node_ref( w, node_child( n, 0, w ));
		nodes[i].child[1] = rc_node_ref( w, node_child( n, 1, w ));
		nodes[i].parent = n->parent ? (int)( n->parent - w->nodes ) : -1;
		nodes[i].firstsurface = node_firstsurface( n, w );
		nodes[i].numsurfaces = node_numsurfaces( n, w );
		memcpy( nodes[i].minmaxs, n->minmaxs, sizeof( n->minmaxs ));
	}

	/* the PVS rows live inside model_t.visdata; a leaf's row is addressed by offset */
	size_t vislen = 0;
	rc_leaf_t *leafs = rc_malloc( sizeof( rc_leaf_t ) * numleafs );
	for( int i = 0; i < numleafs; i++ )
	{
		const mleaf_t *l = &w->leafs[i];
		leafs[i].contents = l->contents;
		leafs[i].cluster = l->cluster;
		leafs[i].parent = l->parent ? (int)( l->parent - w->nodes ) : -1;
		leafs[i].firstmark = l->firstmarksurface ? (int)( l->firstmarksurface - w->marksurfaces ) : 0;
		leafs[i].nummark = l->nummarksurfaces;
		leafs[i].visofs = ( l->compressed_vis && w->visdata ) ? (int)( l->compressed_vis - w->visdata ) : -1;
		memcpy( leafs[i].minmaxs, l->minmaxs, sizeof( l->minmaxs ));
		if( leafs[i].visofs >= 0 )
		{
			/* walk the row to find where the lump ends (model_t does not keep the lump size) */
			const byte *in = l->compressed_vis;
			int d = 0;
			while( d < sc.visbytes )
			{
				if( *in ) { d++; in++; }
				else { d += in[1]; in += 2; }
			}
			size_t end = (size_t)( in - w->visdata );
			if( end > vislen ) vislen = end;
		}
	}
	sc.visdata_len = w->visdata ? (int)Q_max( vislen, 1 ) : 0;

	int32_t *marks = rc_malloc( sizeof( int32_t ) * Q_max( 1, w->nummarksurfaces ));
	for( int i = 0; i < w->nummarksurfaces; i++ )
		marks[i] = (int)( w->marksurfaces[i] - w->surfaces );

	uint16_t *edges = rc_malloc( sizeof( uint16_t ) * 2 * Q_max( 1, w->numedges ));
	for( int i = 0; i < w->numedges; i++ )
	{
		edges[2 * i] = w->edges16[i].v[0];
		edges[2 * i + 1] = w->edges16[i].v[1];
	}

	/* lighting: byte length = end of the furthest sample block */
	size_t lightlen = 0;
	rc_free( rc.surf_texnum );
	rc.surf_texnum = rc_malloc( sizeof( int32_t ) * Q_max( 1, w->numsurfaces ));
	rc_surface_t *surfs = rc_malloc( sizeof( rc_surface_t ) * Q_max( 1, w->numsurfaces ));
	size_t cache_texels = 0;
	int unsupported = 0;
	for( int i = 0; i < w->numsurfaces; i++ )
	{
		const msurface_t *s = &w->surfaces[i];
		const mextrasurf_t *info = s->info;
		rc_surface_t *d = &surfs[i];
		d->plane = (int)( s->plane - w->planes );
		d->flags = s->flags;
		d->firstedge = s->firstedge;
		d->numedges = s->numedges;
		d->texnum = rc.surf_texnum[i] = rc_surface_texture( s );
		d->sample_size = gEngfuncs.Mod_SampleSizeForFace( s );
		d->texflags = s->texinfo->flags;
		for( int j = 0; j < 4; j++ ) d->styles[j] = s->styles[j];
		for( int j = 0; j < 2; j++ )
		{
			d->texturemins[j] = s->texturemins[j];
			d->extents[j] = s->extents[j];
			d->lightmapmins[j] = info->lightmapmins[j];
			d->lightextents[j] = info->lightextents[j];
			for( int k = 0; k < 4; k++ ) d->lmvecs[j][k] = info->lmvecs[j][k];
		}
		d->lightofs = -1;
		if( s->samples && w->lightdata )
		{
			d->lightofs = (int)( s->samples - w->lightdata );
			int smax = info->lightextents[0] / d->sample_size + 1, tmax = info->lightextents[1] / d->sample_size + 1;
			int nstyles = 0;
			while( nstyles < 4 && s->styles[nstyles] != 255 ) nstyles++;
			size_t end = ((size_t)d->lightofs + (size_t)smax * tmax * nstyles ) * 3;
			if( end > lightlen ) lightlen = end;
		}
		/* every (face,mip) block gets a fixed home in the cache arena: 180 GB of HBM removes the need for the
		 * reference's ring allocator and its evictions (D_SCAlloc, ref/soft/r_surf.c:833) */
		const qboolean nocache = d->sample_size != 16 || FBitSet( s->flags, SURF_CONVEYOR );
		if( !FBitSet( s->flags, SURF_DRAWSKY | SURF_DRAWTURB ) && nocache )
			unsupported++;
		for( int m = 0; m < 4; m++ )
		{
			d->cache_ofs[m] = -1;
			if( FBitSet( s->flags, SURF_DRAWSKY | SURF_DRAWTURB ) || nocache )
				continue;
			int cw = info->lightextents[0] >> m, ch = info->lightextents[1] >> m;
			if( cw <= 0 || ch <= 0 ) continue;
			d->cache_ofs[m] = (int32_t)cache_texels;
			cache_texels += ((size_t)cw * ch + 7 ) & ~(size_t)7;
		}
	}
	sc.lightdata_len = (int)lightlen;
	if( unsupported )
		gEngfuncs.Con_Printf( S_WARN "ref_cuda: %d faces use a lightmap sample size other than 16 (R_DrawSurfaceBlock8_Generic, ref/soft/r_surf.c:345) or are conveyors "
			"(r_surf.c:1065): not built yet, they are filled flat\n", unsupported );

	sc.planes = planes; sc.nodes = nodes; sc.leafs = leafs; sc.surfaces = surfs; sc.marksurfaces = marks;
	sc.surfedges = w->surfedges; sc.edges = edges; sc.verts = (const float *)w->vertexes;
	sc.visdata = w->visdata; sc.lightdata = (const uint8_t *)w->lightdata;

	rc.maxdepth = rc_tree_depth( w, w->nodes, 0 );
	sc.maxdepth = rc.maxdepth;

	rc_sync_textures();
	RC_CHECK_DEV( rc_dev_upload_scene( &sc, cache_texels + 64 ));
	rc.scene_valid = true;

	rc_free( planes ); rc_free( nodes ); rc_free( leafs ); rc_free( marks ); rc_free( edges ); rc_free( surfs );
	gEngfuncs.Con_Reportf( "ref_cuda: world %d faces, %d nodes, %d leafs, tree depth %d, surface cache %.1f MB\n",
		w->numsurfaces, w->numnodes, numleafs, rc.maxdepth, cache_texels * 2.0 / 1048576.0 );
}

/* ---------------------------------------------------------------------------------------------- */
/* light styles: CL_RunLightStyles, ref/common/ref_light.c:30-87                                   */

/* the brightness (x 256) a style shows at its own clock: the letter under the clock's tenth of a second, 'a' = 0 .. 'z' = 25
 * in steps of 22; interpolated styles blend it with the next letter (float, truncated as the reference does) */
static int rc_lightstyle_value( const lightstyle_t *st, qboolean blend )
{
	if( !st->length )
		return 256; /* no pattern: normal light */
	if( st->length == 1 )
		return st->map[0] * 22;
	const int tick = (int)Q_floor( st->time * 10 );
	if( !( blend && st->interp ))
		return st->map[tick % st->length] * 22;
	const int next = (int)Q_ceil( st->time * 10 );
	const float towards_next = ( st->time * 10 ) - tick, towards_this = 1.0f - towards_next;
	float value = (float)( st->map[tick % st->length] * 22.0f ) * towards_this;
	value += (float)( st->map[next % st->length] * 22.0f ) * towards_next;
	return (int)value;
}

/* the CL_RunLightStyles slot (ref/common/ref_light.c:30): advance every style's clock by the frame time, then sample all 256 */
static void CL_RunLightStyles( lightstyle_t *ls )
{
	const model_t *world = gp_cl->models[1];
	if( !world )
		return;
	const qboolean everything_lit = r_fullbright->value || !world->lightdata;
	const float dt = gp_cl->time - gp_cl->oldtime;
	const qboolean clocks_run = !everything_lit && !gp_cl->paused && dt <= 0.1f;
	const qboolean blend = cl_lightstyle_lerping->value != 0;
	for( int i = 0; i < MAX_LIGHTSTYLES; i++ )
	{
		if( clocks_run )
			ls[i].time += dt;
		rc.lightstylevalue[i] = everything_lit ? 256 * 256 : rc_lightstyle_value( &ls[i], blend );
	}
}

/* ---------------------------------------------------------------------------------------------- */
/* per-view constants                                                                               */

/*
 * RI.worldviewProjectionMatrix (ref/soft/r_main.c:405-435, 1078-1081; ref/common/ref_math.c): the GL-style projection
 * times the quake-to-GL axis swap, rotated by -roll, -pitch, -yaw and moved by -vieworigin.  Only R_WorldToScreen reads
 * it (tracers, the WorldToScreen slot).  4x4 products with the reference's summation order, rotations by its formula.
 */
typedef float rc_m44_t[4][4];

static void rc_m44_mul( rc_m44_t out, rc_m44_t a, rc_m44_t b )
{
	rc_m44_t r;
	for( int i = 0; i < 4; i++ )
		for( int j = 0; j < 4; j++ )
			r[i][j] = a[i][0] * b[0][j] + a[i][1] * b[1][j] + a[i][2] * b[2][j] + a[i][3] * b[3][j];
	memcpy( out, r, sizeof( r ));
}

/* m = m * rotation by `angle` degrees about the axis (x, y, z) (Matrix4x4_ConcatRotate) */
static void rc_m44_rotate( rc_m44_t m, float angle, float x, float y, float z )
{
	float len = x * x + y * y + z * z;
	if( len != 0.0f )
		len = 1.0f / sqrt( len );
	x *= len; y *= len; z *= len;
	float c, s;
	angle *= ( -M_PI_F / 180.0f );
	SinCos( angle, &s, &c );
	rc_m44_t r = {
		{ x * x + c * ( 1 - x * x ), x * y * ( 1 - c ) + z * s, z * x * ( 1 - c ) - y * s, 0.0f },
		{ x * y * ( 1 - c ) - z * s, y * y + c * ( 1 - y * y ), y * z * ( 1 - c ) + x * s, 0.0f },
		{ z * x * ( 1 - c ) + y * s, y * z * ( 1 - c ) - x * s, z * z + c * ( 1 - z * z ), 0.0f },
		{ 0.0f, 0.0f, 0.0f, 1.0f } };
	rc_m44_mul( m, m, r );
}

/* R_WorldToScreen (r_main.c:210) through rows 0, 1, 3 of a view's matrix; true = behind the eye (x, y left undivided) */
static int rc_world_to_screen_m( const float m[3][4], const float *world, float *screen )
{
	screen[0] = m[0][0] * world[0] + m[0][1] * world[1] + m[0][2] * world[2] + m[0][3];
	screen[1] = m[1][0] * world[0] + m[1][1] * world[1] + m[1][2] * world[2] + m[1][3];
	const float w = m[2][0] * world[0] + m[2][1] * world[1] + m[2][2] * world[2] + m[2][3];
	screen[2] = 0.0f;
	if( w < 0.001f )
		return true;
	const float invw = 1.0f / w;
	screen[0] *= invw;
	screen[1] *= invw;
	return false;
}

static void rc_view_wvp( const ref_viewpass_t *rvp, rc_view_t *v )
{
	/* R_SetupModelviewMatrix: x right / y up / -z forward out of quake's x forward / y left / z up */
	rc_m44_t mv = { { 0.0f, -1.0f, 0.0f, 0.0f }, { 0.0f, 0.0f, 1.0f, 0.0f }, { -1.0f, 0.0f, 0.0f, 0.0f }, { 0.0f, 0.0f, 0.0f, 1.0f } };
	rc_m44_rotate( mv, -rvp->viewangles[2], 1, 0, 0 );
	rc_m44_rotate( mv, -rvp->viewangles[0], 0, 1, 0 );
	rc_m44_rotate( mv, -rvp->viewangles[1], 0, 0, 1 );
	rc_m44_t tr = { { 1.0f, 0.0f, 0.0f, -rvp->vieworigin[0] }, { 0.0f, 1.0f, 0.0f, -rvp->vieworigin[1] },
		{ 0.0f, 0.0f, 1.0f, -rvp->vieworigin[2] }, { 0.0f, 0.0f, 0.0f, 1.0f } };
	rc_m44_mul( mv, mv, tr );

	/* R_SetupProjectionMatrix, R_GetFarClip */
	const movevars_t *mvars = (const movevars_t *)gEngfuncs.EngineGetParm( PARM_GET_MOVEVARS_PTR, 0 );
	const float farclip = ( rc.world && FBitSet( rvp->flags, RF_DRAW_WORLD ) && mvars ) ? mvars->zmax * 1.73f : 2048.0f;
	const float zNear = 4.0f, zFar = Q_max( 256.0f, farclip );
	const float yMax = zNear * tan( rvp->fov_y * M_PI_F / 360.0f ), yMin = -yMax;
	const float xMax = zNear * tan( rvp->fov_x * M_PI_F / 360.0f ), xMin = -xMax;
	rc_m44_t pr;
	memset( pr, 0, sizeof( pr ));
	pr[0][0] = ( 2.0f * zNear ) / ( xMax - xMin );
	pr[1][1] = ( 2.0f * zNear ) / ( yMax - yMin );
	pr[2][2] = -( zFar + zNear ) / ( zFar - zNear );
	pr[0][2] = ( xMax + xMin ) / ( xMax - xMin );
	pr[1][2] = ( yMax + yMin ) / ( yMax - yMin );
	pr[3][2] = -1.0f;
	pr[2][3] = -( 2.0f * zFar * zNear ) / ( zFar - zNear );

	rc_m44_t wvp;
	rc_m44_mul( wvp, pr, mv );
	memcpy( v->wvp[0], wvp[0], sizeof( wvp[0] ));
	memcpy( v->wvp[1], wvp[1], sizeof( wvp[1] ));
	memcpy( v->wvp[2], wvp[3], sizeof( wvp[3] ));
}

/*
 * R_SetupFrustum (ref/soft/r_main.c:387) + R_SetupFrameQ / R_ViewChanged / R_TransformFrustum /
 * R_SetUpFrustumIndexes (r_misc.c:101-322) for one ref_viewpass_t, with the reference's float
 * expression order and this host's libm (tgmath: tan on a float argument is tanf).
 */
static void rc_setup_view( const ref_viewpass_t *rvp, rc_view_t *v )
{
	vec3_t vforward, vright, vup;
	memset( v, 0, sizeof( *v ));

	AngleVectors( rvp->viewangles, vforward, vright, vup );
	VectorCopy( rvp->vieworigin, v->origin );
	VectorCopy( vforward, v->vforward ); VectorCopy( vright, v->vright ); VectorCopy( vup, v->vup );

	int x = rvp->viewport[0], y = rvp->viewport[1], w = rvp->viewport[2], h = rvp->viewport[3];
	v->vrect[0] = x; v->vrect[1] = y; v->vrect[2] = w; v->vrect[3] = h;

	float horizontalFieldOfView = 2 * tan((float)rvp->fov_x / 360.0f * M_PI_F );
	float verticalFieldOfView = 2 * tan((float)rvp->fov_y / 360.0f * M_PI_F );

	int vrectright = x + w, vrectbottom = y + h;
	v->fvrectx_adj = (float)x - 0.5f;
	v->vrect_x_adj_shift20 = ( x << 20 ) + ( 1 << 19 ) - 1;
	v->fvrecty_adj = (float)y - 0.5f;
	v->vrectright_adj_shift20 = ( vrectright << 20 ) + ( 1 << 19 ) - 1;
	v->fvrectright_adj = (float)vrectright - 0.5f;
	v->fvrectbottom_adj = (float)vrectbottom - 0.5f;

	float xOrigin = 0.5f, yOrigin = 0.5f; /* XCENTERING, YCENTERING */
	v->xcenter = ((float)w * 0.5f ) + x - 0.5f;
	v->ycenter = ((float)h * 0.5f ) + y - 0.5f;
	v->xscale = w / horizontalFieldOfView;
	v->xscaleinv = 1.0f / v->xscale;
	v->yscale = v->xscale;
	v->yscaleinv = 1.0f / v->yscale;

	vec3_t screenedge[4];
	screenedge[0][0] = -1.0f / ( xOrigin * horizontalFieldOfView ); screenedge[0][1] = 0; screenedge[0][2] = 1;
	screenedge[1][0] = 1.0f / (( 1.0f - xOrigin ) * horizontalFieldOfView ); screenedge[1][1] = 0; screenedge[1][2] = 1;
	screenedge[2][0] = 0; screenedge[2][1] = -1.0f / ( yOrigin * verticalFieldOfView ); screenedge[2][2] = 1;
	screenedge[3][0] = 0; screenedge[3][1] = 1.0f / (( 1.0f - yOrigin ) * verticalFieldOfView ); screenedge[3][2] = 1;
	for( int i = 0; i < 4; i++ )
		VectorNormalize( screenedge[i] );
	for( int i = 0; i < 4; i++ )
		VectorCopy( screenedge[i], v->screenedge[i] );

	/* R_TransformFrustum */
	for( int i = 0; i < 4; i++ )
	{
		vec3_t t, v2;
		t[0] = screenedge[i][2];
		t[1] = -screenedge[i][0];
		t[2] = screenedge[i][1];
		v2[0] = t[1] * vright[0] + t[2] * vup[0] + t[0] * vforward[0];
		v2[1] = t[1] * vright[1] + t[2] * vup[1] + t[0] * vforward[1];
		v2[2] = t[1] * vright[2] + t[2] * vup[2] + t[0] * vforward[2];
		VectorCopy( v2, v->clip_normal[i] );
		v->clip_dist[i] = DotProduct( rvp->vieworigin, v2 );
	}
	/* R_SetUpFrustumIndexes */
	for( int i = 0; i < 4; i++ )
		for( int j = 0; j < 3; j++ )
		{
			if( v->clip_normal[i][j] < 0 ) { v->frustum_idx[i][j] = j; v->frustum_idx[i][j + 3] = j + 3; }
			else { v->frustum_idx[i][j] = j + 3; v->frustum_idx[i][j + 3] = j; }
		}

	/* D_ViewChanged + d_setup */
	v->scale_for_mip = v->xscale;
	if( v->yscale > v->xscale ) v->scale_for_mip = v->yscale;
	int minmip = sw_mipcap.value;
	v->minmip = minmip > 3 ? 3 : ( minmip < 0 ? 0 : minmip );
	static const float basemip[3] = { 1.0, 0.5 * 0.8, 0.25 * 0.8 };
	for( int i = 0; i < 3; i++ )
		v->scalemip[i] = basemip[i] * sw_mipscale.value;

	/* D_DrawSurfaces: transformed_modelorg = TransformVector( vieworigin - vec3_origin ) */
	vec3_t modelorg;
	VectorSubtract( rvp->vieworigin, vec3_origin, modelorg );
	v->modelorg_t[0] = DotProduct( modelorg, vright );
	v->modelorg_t[1] = DotProduct( modelorg, vup );
	v->modelorg_t[2] = DotProduct( modelorg, vforward );

	/* R_SetUpWorldTransform (ref/soft/r_trialias.c:52): the view matrix (vright, -vup, vforward) times a translation by
	 * -vieworigin, through the engine's own Matrix3x4_ConcatTransforms so that every product rounds as it does there */
	{
		matrix3x4 viewmatrix, world;
		vec3_t zero = { 0, 0, 0 }, f, r, u;
		AngleVectors( zero, f, r, u );
		memset( world, 0, sizeof( world ));
		for( int i = 0; i < 3; i++ ) { world[i][0] = f[i]; world[i][1] = -r[i]; world[i][2] = u[i]; }
		world[0][3] = -rvp->vieworigin[0]; world[1][3] = -rvp->vieworigin[1]; world[2][3] = -rvp->vieworigin[2];
		VectorCopy( vright, viewmatrix[0] );
		VectorCopy( vup, viewmatrix[1] );
		viewmatrix[1][0] = -viewmatrix[1][0]; viewmatrix[1][1] = -viewmatrix[1][1]; viewmatrix[1][2] = -viewmatrix[1][2];
		VectorCopy( vforward, viewmatrix[2] );
		viewmatrix[0][3] = viewmatrix[1][3] = viewmatrix[2][3] = 0;
		matrix3x4 out;
		Matrix3x4_ConcatTransforms( out, viewmatrix, world );
		memcpy( v->aliastransform, out, sizeof( out ));
	}

	v->turb_phase = (int)( gp_cl->time * 20 ) & 127; /* SPEED, CYCLE */
	v->sky_color = (int)sw_clearcolor.value & 0xFFFF;
	v->texfilt = sw_texfilt.value == 1.0f;
	v->flags = rvp->flags;
	v->pvs_radius = r_novis.value ? -1.0f : r_pvs_radius->value;
	/* R_WorldToScreen's matrix: three rotations and two tangents per view, only for batches that ask for it */
	if( rc.batch_wvp )
		rc_view_wvp( rvp, v );
	rc.last_rvp = *rvp;
}

/* ---------------------------------------------------------------------------------------------- */
/* frames: views queue up between R_BeginFrame and R_EndFrame and are rendered as one batch          */

/* R_PushDlights (ref/common/ref_light.c:170): the engine's dynamic lights that are alive at the frame's time; the
 * faces each one reaches are marked on the device (k_dlight_mark), the light is added by k_cache_build */
static void rc_push_dlights( void )
{
	const dlight_t *dl = (const dlight_t *)gEngfuncs.EngineGetParm( PARM_GET_DLIGHTS_PTR, 0 );
	rc_dlight_t act[RC_MAX_DLIGHTS];
	int n = 0;
	for( int i = 0; dl && i < MAX_DLIGHTS && i < RC_MAX_DLIGHTS; i++ )
	{
		const dlight_t *l = &dl[i];
		if( l->die < gp_cl->time || !l->radius )
			continue;
		rc_dlight_t *o = &act[n++];
		VectorCopy( l->origin, o->origin );
		o->radius = l->radius;
		o->minlight = l->minlight;
		uint b = ( l->color.r + l->color.g + l->color.b ) / 3 * 4;
		o->monolight = (int)(( b >= 1024 ? 0u : (uint)rc.lightgamma[b] ) * 3 );
		o->bit = i;
		o->virtual_radius = l->radius * Q_max( 1.0f, r_dlight_virtual_radius.value );
	}
	if( n || rc.had_dlights )
		RC_CHECK_DEV( rc_dev_set_dlights( act, n ));
	rc.had_dlights = n > 0;
}


static const rc_texture_t *rc_texture_entry( int texnum )
{
	return ( texnum > 0 && texnum < rc.ntextab && RC_IMAGE_LIVE( &rc.images[texnum] )) ? &rc.textab[texnum] : NULL;
}

/* ---------------------------------------------------------------------------------------------- */
/* blend tables: vid.addmap / modmap / alphamap of R_BuildBlendMaps (ref/soft/r_glblit.c:411-520).  A pixel_t is a
 * 3-3-2 "major" byte over a byte of low channel bits; addmap and modmap act on major bytes (saturating add, scaled
 * product), alphamap mixes a 10-bit screen code (major + 2 minor bits) with a major source code at 2/7, 3/7, 4/7. */

static void rc_build_blendmaps( uint16_t *alphamap, uint8_t *addmap, uint8_t *modmap )
{
	for( uint a = 0; a < 256; a++ )
		for( uint b = 0; b < 256; b++ )
		{
			const uint r1 = a >> 5, g1 = ( a >> 2 ) & 7, b1 = a & 3, r2 = b >> 5, g2 = ( b >> 2 ) & 7, b2 = b & 3;
			addmap[b << 8 | a] = (uint8_t)( Q_min( r1 + r2, 7u ) << 5 | Q_min( g1 + g2, 7u ) << 2 | Q_min( b1 + b2, 3u ));
			modmap[b << 8 | a] = (uint8_t)(( r1 * r2 / 7 ) << 5 | ( g1 * g2 / 7 ) << 2 | ( b1 * b2 / 3 ));
		}
	for( uint i = 0; i < 1024; i++ )
	{
		uint r1, g1, b1;
		rc_decode565( i << 6 | 0x38, &r1, &g1, &b1 ); /* the screen colour: 10 known bits, the three below them set */
		for( uint m2 = 0; m2 < 256; m2++ )
		{
			const uint r2 = ( m2 >> 5 ) << 2 | 4, g2 = (( m2 >> 2 ) & 7 ) << 3 | 3, b2 = ( m2 & 3 ) << 3 | 3; /* the source's major byte, mid-range */
			for( uint k = 0; k < 3; k++ )
			{
				const uint a = k + 2;
				const uint r = Q_min( r1 * ( 7 - a ) / 7 + r2 * a / 7, 31u ), g = Q_min( g1 * ( 7 - a ) / 7 + g2 * a / 7, 63u ), b = Q_min( b1 * ( 7 - a ) / 7 + b2 * a / 7, 31u );
				alphamap[k << 18 | m2 << 10 | i] = rc_encode565( r, g, b ) & ~0x3f;
			}
		}
	}
}

/* ---------------------------------------------------------------------------------------------- */
/* entity lists: R_AddEntity / R_ClearScene (ref/soft/r_main.c:293-363), CL_FxBlend (:1406)           */

static int rc_fx_blend( cl_entity_t *e )
{
	int blend = 0;
	const float offset = ((int)e->index ) * 363.0f;
	const int amt = e->curstate.renderamt;
	switch( e->curstate.renderfx )
	{
	case kRenderFxPulseSlowWide: blend = amt + 0x40 * sin( gp_cl->time * 2 + offset ); break;
	case kRenderFxPulseFastWide: blend = amt + 0x40 * sin( gp_cl->time * 8 + offset ); break;
	case kRenderFxPulseSlow: blend = amt + 0x10 * sin( gp_cl->time * 2 + offset ); break;
	case kRenderFxPulseFast: blend = amt + 0x10 * sin( gp_cl->time * 8 + offset ); break;
	case kRenderFxFadeSlow: e->curstate.renderamt = amt > 0 ? amt - 1 : 0; blend = e->curstate.renderamt; break;
	case kRenderFxFadeFast: e->curstate.renderamt = amt > 3 ? amt - 4 : 0; blend = e->curstate.renderamt; break;
	case kRenderFxSolidSlow: e->curstate.renderamt = amt < 255 ? amt + 1 : 255; blend = e->curstate.renderamt; break;
	case kRenderFxSolidFast: e->curstate.renderamt = amt < 252 ? amt + 4 : 255; blend = e->curstate.renderamt; break;
	case kRenderFxStrobeSlow: blend = 20 * sin( gp_cl->time * 4 + offset ); blend = blend < 0 ? 0 : amt; break;
	case kRenderFxStrobeFast: blend = 20 * sin( gp_cl->time * 16 + offset ); blend = blend < 0 ? 0 : amt; break;
	case kRenderFxStrobeFaster: blend = 20 * sin( gp_cl->time * 36 + offset ); blend = blend < 0 ? 0 : amt; break;
	case kRenderFxFlickerSlow: blend = 20 * ( sin( gp_cl->time * 2 ) + sin( gp_cl->time * 17 + offset )); blend = blend < 0 ? 0 : amt; break;
	case kRenderFxFlickerFast: blend = 20 * ( sin( gp_cl->time * 16 ) + sin( gp_cl->time * 23 + offset )); blend = blend < 0 ? 0 : amt; break;
	case kRenderFxHologram:
	case kRenderFxDistort:
		/* the reference fades these with the distance along the CURRENT view's forward axis; a batch has many views */
		if( !rc.warned[0] ) { rc.warned[0] = true; gEngfuncs.Con_Reportf( "ref_cuda: kRenderFxHologram / kRenderFxDistort fade per view is not built: drawn at renderamt 180\n" ); }
		e->curstate.renderamt = 180;
		blend = 180;
		break;
	default: blend = amt; break;
	}
	return bound( 0, blend, 255 );
}

static qboolean rc_model_opaque( int rendermode ) { return rendermode == kRenderNormal; }

/* R_GetEntityRenderMode (ref/soft/r_studio.c:2109) */
static int rc_entity_rendermode( cl_entity_t *ent )
{
	model_t *model = ent->model;
	studiohdr_t *phdr = model ? (studiohdr_t *)gEngfuncs.Mod_Extradata( mod_studio, model ) : NULL;
	if( !phdr || !model || model->type != mod_studio )
	{
		if( rc_model_opaque( ent->curstate.rendermode ) && model && model->type == mod_brush && FBitSet( model->flags, MODEL_TRANSPARENT ))
			return kRenderTransAlpha;
		return ent->curstate.rendermode;
	}
	const mstudiotexture_t *ptexture = (const mstudiotexture_t *)((const byte *)phdr + phdr->textureindex );
	int opaque = 0, trans = 0;
	for( int i = 0; i < phdr->numtextures; i++, ptexture++ )
	{
		if( FBitSet( ptexture->flags, STUDIO_NF_ADDITIVE ) && !FBitSet( ptexture->flags, STUDIO_NF_CHROME )) trans++;
		else opaque++;
	}
	return trans > opaque ? kRenderTransAdd : ent->curstate.rendermode;
}

static qboolean rc_opaque_entity( cl_entity_t *ent )
{
	if( rc_entity_rendermode( ent ) == kRenderNormal )
	{
		switch( ent->curstate.renderfx )
		{
		case kRenderFxNone: case kRenderFxDeadPlayer: case kRenderFxLightMultiplier: case kRenderFxExplode:
			return true;
		}
	}
	return false;
}

static qboolean R_AddEntity( struct cl_entity_s *clent, int type )
{
	if( !r_drawentities->value )
		return false;
	if( FBitSet( clent->curstate.effects, EF_NODRAW ))
		return false;
	if( !rc_model_opaque( clent->curstate.rendermode ) && rc_fx_blend( clent ) <= 0 )
		return true; /* invisible */
	if( type == ET_BEAM )
	{
		if( !rc.warned[1] ) { rc.warned[1] = true; gEngfuncs.Con_Reportf( "ref_cuda: beam entities are not drawn yet\n" ); }
		return true;
	}
	if( rc_opaque_entity( clent ))
	{
		if( clent->model && clent->model->type == mod_brush )
		{
			if( rc.num_edge >= RC_MAX_VISIBLE ) return false;
			rc.edge_ents[rc.num_edge++] = clent;
			return true;
		}
		if( rc.num_solid >= RC_MAX_VISIBLE ) return false;
		rc.solid_ents[rc.num_solid++] = clent;
	}
	else
	{
		if( rc.num_trans >= RC_MAX_VISIBLE ) return false;
		rc.trans_ents[rc.num_trans++] = clent;
	}
	return true;
}

static void R_ClearScene( void ) { rc.num_solid = rc.num_trans = rc.num_edge = 0; }
static void R_ProcessEntData( qboolean allocate, cl_entity_t *entities, unsigned int max_entities )
{
	rc.entities = allocate ? entities : NULL;
	rc.max_entities = allocate ? max_entities : 0;
}

/* ---------------------------------------------------------------------------------------------- */
/* TriAPI (ref/soft/r_triapi.c:46-261) and GL_Bind (r_image.c:43-75): the calls keep the reference's immediate-mode
 * state; where the reference rasterises a triangle on the spot, here its three world-space corners join the batch's
 * draw items - projection, clipping and rasterisation happen per view on the device.                                  */

static void *rc_grow( void *p, int *cap, int need, size_t elem )
{
	if( need <= *cap ) return p;
	int n = *cap ? *cap : 256;
	while( n < need ) n *= 2;
	p = gEngfuncs._Mem_Realloc( rc.pool, p, elem * n, true, __FILE__, __LINE__ );
	*cap = n;
	return p;
}

static rc_item_t *rc_new_item( int kind )
{
	rc.items = rc_grow( rc.items, &rc.maxitems, rc.nitems + 1, sizeof( rc_item_t ));
	rc_item_t *it = &rc.items[rc.nitems++];
	memset( it, 0, sizeof( *it ));
	it->kind = kind; it->part = -1; it->glow = -1;
	it->drawer = rc.tri.drawer; it->tex_ofs = rc.tri.texnum; /* texture NUMBER here; resolved to an arena offset when the batch is sent */
	it->skinwidth = rc.tri.skinwidth; it->skinheight = rc.tri.skinheight;
	it->color = rc.tri.color; it->alpha = rc.tri.alpha;
	it->sortent = rc.cur_sortent;
	return it;
}

static void GL_SetRenderMode( int renderMode ) { rc.tri.rendermode = renderMode; }
static void TriRenderMode( int mode ) { rc.tri.triRenderMode = rc.tri.rendermode = mode; }

static void GL_Bind( int tmu, unsigned int texnum )
{
	(void)tmu;
	if( texnum >= RC_MAX_TEXTURES ) texnum = 0;
	const rc_image_t *image = &rc.images[texnum];
	const int mode = rc.tri.rendermode;
	rc.tri.drawer = mode == kRenderNormal ? RC_DRAW_FILL : mode == kRenderTransAdd ? RC_DRAW_ADDITIVE : mode == kRenderGlow ? RC_DRAW_GLOW
		: image->has_alpha ? RC_DRAW_TEXBLEND : RC_DRAW_BLEND;
	rc.tri.texnum = (int)texnum;
	rc.tri.skinwidth = image->width; rc.tri.skinheight = image->height;
}

static void rc_tri_color4f( float rr, float gg, float bb, float aa )
{
	if( rc.tri.rendermode == kRenderTransAdd || rc.tri.rendermode == kRenderGlow )
		rr *= aa, gg *= aa, bb *= aa;
	rc.tri.light = ( rr + gg + bb ) * 31 / 3;
	if( rc.tri.light > 31 ) rc.tri.light = 31;
	if( !rc.tri.is2d && rc.tri.rendermode == kRenderNormal )
		return;
	rc.tri.alpha = aa * 7;
	if( rc.tri.alpha > 7 ) rc.tri.alpha = 7;
	if( rr == 1 && gg == 1 && bb == 1 )
	{
		rc.tri.color = 0xFFFF; /* COLOR_WHITE */
		return;
	}
	unsigned short r = rr * 31, g = gg * 63, b = bb * 31;
	rc.tri.color = rc_encode565( Q_min( r, 31 ), Q_min( g, 63 ), Q_min( b, 31 ));
}

/* _TriColor4ub (ref/soft/r_triapi.c:143) */
static void rc_tri_color4ub( unsigned char r, unsigned char g, unsigned char b, unsigned char a )
{
	rc_tri_color4f( r * ( 1.0f / 255.0f ), g * ( 1.0f / 255.0f ), b * ( 1.0f / 255.0f ), a * ( 1.0f / 255.0f ));
}

static void TriColor4ub( unsigned char r, unsigned char g, unsigned char b, unsigned char a )
{
	rc.tri.triRGBA[0] = r * ( 1.0f / 255.0f ); rc.tri.triRGBA[1] = g * ( 1.0f / 255.0f );
	rc.tri.triRGBA[2] = b * ( 1.0f / 255.0f ); rc.tri.triRGBA[3] = a * ( 1.0f / 255.0f );
	rc_tri_color4f( rc.tri.triRGBA[0], rc.tri.triRGBA[1], rc.tri.triRGBA[2], 1.0f );
}

static void TriColor4f( float r, float g, float b, float a )
{
	if( rc.tri.triRenderMode == kRenderTransAlpha )
		TriColor4ub( r * 255.0f, g * 255.0f, b * 255.0f, a * 255.0f );
	else
		rc_tri_color4f( r * a, g * a, b * a, 1.0 );
	rc.tri.triRGBA[0] = r; rc.tri.triRGBA[1] = g; rc.tri.triRGBA[2] = b; rc.tri.triRGBA[3] = a;
}

static void TriTexCoord2f( float u, float v )
{
	double u1 = 0, v1 = 0;
	u = fmodf( u, 10 );
	v = fmodf( v, 10 );
	if( u < 1000 && u > -1000 ) u1 = u;
	if( v < 1000 && v > -1000 ) v1 = v;
	while( u1 < 0 ) u1 = u1 + 1;
	while( v1 < 0 ) v1 = v1 + 1;
	while( u1 > 1 ) u1 = u1 - 1;
	while( v1 > 1 ) v1 = v1 - 1;
	rc.tri.s = rc.tri.skinwidth * bound( 0.01, u1, 0.99 );
	rc.tri.t = rc.tri.skinheight * bound( 0.01, v1, 0.99 );
}

static void TriBegin( int primitiveCode )
{
	if( primitiveCode == TRI_QUADS ) primitiveCode = TRI_TRIANGLE_FAN;
	rc.tri.prim = primitiveCode;
	rc.tri.n = rc.tri.vertcount = 0;
}
static void TriEnd( void ) {}

/* one assembled triangle (R_RenderTriangle's arguments) joins the current item, or opens one when the state changed */
static void rc_tri_emit( const rc_corner_t *a, const rc_corner_t *b, const rc_corner_t *c )
{
	rc_item_t *it = rc.nitems ? &rc.items[rc.nitems - 1] : NULL;
	if( !it || it->kind != RC_ITEM_TRIS || it->drawer != rc.tri.drawer || it->tex_ofs != rc.tri.texnum || it->color != (int)rc.tri.color
		|| it->alpha != (int)rc.tri.alpha || it->sortent != rc.cur_sortent || it->first + 3 * it->count != rc.ncorners )
	{
		it = rc_new_item( RC_ITEM_TRIS );
		it->first = rc.ncorners;
	}
	rc.corners = rc_grow( rc.corners, &rc.maxcorners, rc.ncorners + 3, sizeof( rc_corner_t ));
	rc.corners[rc.ncorners++] = *a; rc.corners[rc.ncorners++] = *b; rc.corners[rc.ncorners++] = *c;
	it->count++;
}

static void TriVertex3f( float x, float y, float z )
{
	rc_corner_t cv;
	cv.xyz[0] = x; cv.xyz[1] = y; cv.xyz[2] = z;
	cv.l = (int32_t)( rc.tri.light << 8 ); cv.s = rc.tri.s; cv.t = rc.tri.t;
	rc_corner_t *triv = rc.tri.triv;
	if( rc.tri.prim == TRI_TRIANGLES )
	{
		triv[rc.tri.vertcount++] = cv;
		if( rc.tri.vertcount == 3 )
		{
			rc_tri_emit( &triv[0], &triv[1], &triv[2] );
			rc.tri.vertcount = 0;
		}
	}
	if( rc.tri.prim == TRI_TRIANGLE_FAN )
	{
		triv[rc.tri.vertcount++] = cv;
		if( rc.tri.vertcount >= 3 )
		{
			rc_tri_emit( &triv[0], &triv[1], &triv[2] );
			triv[1] = triv[2];
			rc.tri.vertcount = 2;
		}
	}
	if( rc.tri.prim == TRI_TRIANGLE_STRIP )
	{
		triv[rc.tri.n] = cv;
		rc.tri.n++;
		rc.tri.vertcount++;
		if( rc.tri.n == 3 ) rc.tri.n = 0;
		if( rc.tri.vertcount >= 3 )
		{
			if( rc.tri.vertcount & 1 ) rc_tri_emit( &triv[0], &triv[1], &triv[2] );
			else rc_tri_emit( &triv[2], &triv[1], &triv[0] );
		}
	}
}
static void TriVertex3fv( const float *worldPnt ) { TriVertex3f( worldPnt[0], worldPnt[1], worldPnt[2] ); }
static void TriCullFace( TRICULLSTYLE mode ) { (void)mode; }

/* ---------------------------------------------------------------------------------------------- */
/* sprites: the view-independent half of R_DrawSpriteModel (ref/soft/r_sprite.c:160-266); orientation, the quad's four
 * corners and everything after them run per view on the device                                                       */

static void rc_add_sprite( cl_entity_t *e )
{
	model_t *model = e->model;
	msprite_t *psprite = (msprite_t *)model->cache.data;
	if( !psprite )
		return; /* R_CullSpriteModel */
	vec3_t origin;
	VectorCopy( e->origin, origin );
	if( e->curstate.aiment > 0 && e->curstate.movetype == MOVETYPE_FOLLOW && rc.entities && (uint)e->curstate.aiment < rc.max_entities )
	{
		cl_entity_t *parent = &rc.entities[e->curstate.aiment];
		if( parent->model )
		{
			if( parent->model->type == mod_studio && e->curstate.body > 0 )
				VectorCopy( parent->attachment[bound( 1, e->curstate.body, MAXSTUDIOATTACHMENTS ) - 1], origin );
			else VectorCopy( parent->origin, origin );
		}
	}
	float scale = e->curstate.scale;
	if( !scale ) scale = 1.0f;
	GL_SetRenderMode( e->curstate.rendermode );
	vec3_t color;
	if( e->curstate.rendercolor.r || e->curstate.rendercolor.g || e->curstate.rendercolor.b )
	{
		color[0] = (float)e->curstate.rendercolor.r * ( 1.0f / 255.0f );
		color[1] = (float)e->curstate.rendercolor.g * ( 1.0f / 255.0f );
		color[2] = (float)e->curstate.rendercolor.b * ( 1.0f / 255.0f );
	}
	else color[0] = color[1] = color[2] = 1.0f;
	mspriteframe_t *frame = gEngfuncs.R_GetSpriteFrame( model, e->curstate.frame, e->angles[YAW] );
	if( !frame )
		return;
	int type = psprite->type;
	if( e->angles[ROLL] != 0.0f && type == SPR_FWD_PARALLEL )
		type = SPR_FWD_PARALLEL_ORIENTED;
	rc_tri_color4f( color[0], color[1], color[2], rc.tri.blend );
	GL_Bind( XASH_TEXTURE0, frame->gl_texturenum );

	rc_item_t *it = rc_new_item( RC_ITEM_SPRITE );
	it->count = 2;
	it->sprtype = type; it->light = (int32_t)( rc.tri.light << 8 );
	if( e->curstate.rendermode == kRenderGlow )
	{
		/* R_SpriteOccluded: drawn or not, how large and how bright is decided per view (rc_build_glow, once the batch's
		 * views are known) */
		rc.glowsrc = rc_grow( rc.glowsrc, &rc.maxglow, rc.nglow + 1, sizeof( *rc.glowsrc ));
		struct rc_glowsrc_s *g = &rc.glowsrc[rc.nglow];
		VectorCopy( origin, g->origin ); VectorCopy( color, g->color );
		g->scale = scale; g->blend = rc.tri.blend; g->renderfx = e->curstate.renderfx;
		it->glow = rc.nglow++;
	}
	VectorCopy( origin, it->origin ); VectorCopy( e->angles, it->angles );
	it->scale = scale;
	it->fr_up = frame->up; it->fr_down = frame->down; it->fr_left = frame->left; it->fr_right = frame->right;
	TriTexCoord2f( 0.0f, 0.0f ); it->s0 = rc.tri.s; it->t0 = rc.tri.t;
	TriTexCoord2f( 1.0f, 1.0f ); it->s1 = rc.tri.s; it->t1 = rc.tri.t;
	if( type == SPR_ORIENTED )
	{
		vec3_t v_forward, v_right, v_up;
		AngleVectors( e->angles, v_forward, v_right, v_up );
		VectorScale( v_forward, 0.01f, v_forward ); /* to avoid z-fighting */
		VectorSubtract( origin, v_forward, it->origin );
		VectorCopy( v_right, it->vright ); VectorCopy( v_up, it->vup );
	}
	else if( type == SPR_FWD_PARALLEL_ORIENTED )
	{
		const float angle = e->angles[ROLL] * ( M_PI2 / 360.0f );
		SinCos( angle, &it->sr, &it->cr );
	}
}

static void rc_add_studio( cl_entity_t *e );

/* R_DrawEntitiesOnList (ref/soft/r_main.c:504-635) as a list of draw items: solid studio models, solid sprites, the
 * client's opaque triangles, then the translucent entities (each with a sort record) and the client's translucent ones */
static void rc_add_tbrush( cl_entity_t *e );

static void rc_build_items( void )
{
	rc.nitems = rc.ncorners = rc.nsortents = rc.nglow = rc.ntracers = 0;
	rc.cur_sortent = -1;
	rs.nparts = rs.nbones = rs.totalverts = rs.totalnorms = 0;
	rc.tri.blend = 1.0f;
	rc.tri.drawer = RC_DRAW_FILL;
	GL_SetRenderMode( kRenderNormal );
	for( int i = 0; i < rc.num_solid; i++ )
	{
		cl_entity_t *e = rc.solid_ents[i];
		if( e->model && e->model->type == mod_studio )
			rc_add_studio( e );
	}
	for( int i = 0; i < rc.num_solid; i++ )
	{
		cl_entity_t *e = rc.solid_ents[i];
		if( e->model && e->model->type == mod_sprite )
			rc_add_sprite( e );
	}
	gEngfuncs.CL_DrawEFX( gp_cl->time - gp_cl->oldtime, false );
	gEngfuncs.pfnDrawNormalTriangles();
	for( int i = 0; i < rc.num_trans; i++ )
	{
		cl_entity_t *e = rc.trans_ents[i];
		if( !e->model )
			continue;
		rc.tri.blend = e->curstate.rendermode != kRenderNormal ? rc_fx_blend( e ) / 255.0f : 1.0f;
		if( rc.tri.blend <= 0.0f )
			continue;
		/* R_TransEntityCompare's key for this entity */
		rc.sortents = rc_grow( rc.sortents, &rc.maxsortents, rc.nsortents + 1, sizeof( rc_sortent_t ));
		rc_sortent_t *se = &rc.sortents[rc.nsortents];
		memset( se, 0, sizeof( *se ));
		const int rendermode = rc_entity_rendermode( e );
		vec3_t org;
		VectorAverage( e->model->mins, e->model->maxs, org );
		VectorAdd( e->origin, org, se->center );
		se->use_dist = ( e->model->type != mod_brush || rendermode != kRenderTransAlpha ) ? 1 : 0;
		se->rank = rendermode == kRenderTransTexture ? 1 : rendermode == kRenderTransAdd ? 2 : rendermode == kRenderGlow ? 3 : 0;
		se->first_item = rc.nitems;
		rc.cur_sortent = rc.nsortents++;
		switch( e->model->type )
		{
		case mod_studio: rc_add_studio( e ); break;
		case mod_sprite: rc_add_sprite( e ); break;
		case mod_brush: rc_add_tbrush( e ); break;
		default: break;
		}
		rc.sortents[rc.cur_sortent].nitems = rc.nitems - rc.sortents[rc.cur_sortent].first_item;
	}
	rc.cur_sortent = -1;
	rc.tri.blend = 1.0f;
	gEngfuncs.pfnDrawTransparentTriangles();
	gEngfuncs.CL_DrawEFX( gp_cl->time - gp_cl->oldtime, true );
	GL_SetRenderMode( kRenderNormal );

	/* texture numbers -> texel arena offsets, triangle numbering */
	int order = 0;
	for( int i = 0; i < rc.nitems; i++ )
	{
		rc_item_t *it = &rc.items[i];
		const rc_image_t *img = &rc.images[it->tex_ofs];
		if( it->kind == RC_ITEM_BRUSH ) { it->tex_ofs = 0; it->order0 = order; continue; }
		const rc_texture_t *t = rc_texture_entry( it->tex_ofs );
		it->tex_ofs = t ? ( it->drawer == RC_DRAW_TEXBLEND ? t->alpha_ofs : t->ofs[0] ) : -1;
		(void)img;
		it->order0 = order;
		order += it->count;
	}
	if( rs.dirty )
	{
		RC_CHECK_DEV( rc_dev_upload_studio_static( rs.arena, rs.used ));
		rs.dirty = false;
	}
	RC_CHECK_DEV( rc_dev_set_studio( rs.parts, rs.nparts, rs.bones, rs.nbones, rs.blight, rs.totalverts, rs.totalnorms ));
	RC_CHECK_DEV( rc_dev_set_items( rc.items, rc.nitems, rc.corners, rc.ncorners, rc.sortents, rc.nsortents ));
}

/* ---------------------------------------------------------------------------------------------- */
/* studio models: the per-entity, view-independent half of R_StudioDrawModel (ref/soft/r_studio.c:2720) on the host -
 * skeleton (quaternion math through the engine's `public` library, as ref_soft itself does), entity lighting
 * (R_EntityDynamicLight / R_LightVec, ref/common/ref_light.c:222-700) - and static per-submodel tables for the device,
 * which transforms the vertices by their bones, lights the normals and rasterises the meshes per view.               */


static int32_t rs_put( const void *src, size_t bytes, size_t align )
{
	size_t at = ( rs.used + align - 1 ) & ~( align - 1 );
	if( at + bytes > rs.cap )
	{
		size_t n = rs.cap ? rs.cap : 1 << 20;
		while( n < at + bytes ) n *= 2;
		rs.arena = gEngfuncs._Mem_Realloc( rc.pool, rs.arena, n, true, __FILE__, __LINE__ );
		rs.cap = n;
	}
	memcpy( rs.arena + at, src, bytes );
	rs.used = at + bytes;
	rs.dirty = true;
	return (int32_t)at;
}

/* the static tables of one submodel with one skin family: vertices, normals, and every mesh's strips / fans unrolled into
 * triangles the way TriVertex3f assembles them (r_triapi.c:221), texels from TriTexCoord2f on the mesh's skin */
static int rs_submodel( studiohdr_t *hdr, mstudiomodel_t *sub, int skinnum )
{
	for( int i = 0; i < rs.nsub; i++ )
		if( rs.sub[i].hdr == hdr && rs.sub[i].sub == sub && rs.sub[i].skin == skinnum )
			return i;
	if( rs.nsub == 256 || sub->nummesh > MAXSTUDIOMESHES )
		return -1;
	const int si = rs.nsub++;
	rs.sub[si].hdr = hdr; rs.sub[si].sub = sub; rs.sub[si].skin = skinnum;
	const byte *base = (const byte *)hdr;
	rs.sub[si].sverts = rs_put( base + sub->vertindex, sizeof( vec3_t ) * sub->numverts, 4 ) / 4;
	rs.sub[si].snorms = rs_put( base + sub->normindex, sizeof( vec3_t ) * sub->numnorms, 4 ) / 4;
	rs.sub[si].svertbone = rs_put( base + sub->vertinfoindex, sub->numverts, 4 );
	rs.sub[si].snormbone = rs_put( base + sub->norminfoindex, sub->numnorms, 4 );
	const mstudiotexture_t *ptexture = (const mstudiotexture_t *)( base + hdr->textureindex );
	const short *pskinref = (const short *)( base + hdr->skinindex );
	if( skinnum > 0 && skinnum < hdr->numskinfamilies )
		pskinref += skinnum * hdr->numskinref;
	const mstudiomesh_t *pmesh = (const mstudiomesh_t *)( base + sub->meshindex );
	byte *nflags = rc_malloc( Q_max( 1, sub->numnorms ));
	int k = 0;
	for( int j = 0; j < sub->nummesh; j++ )
		for( int i = 0; i < pmesh[j].numnorms && k < sub->numnorms; i++, k++ )
			nflags[k] = (byte)( ptexture[pskinref[pmesh[j].skinref]].flags & 0xFF );
	rs.sub[si].snormflags = rs_put( nflags, Q_max( 1, sub->numnorms ), 4 );
	rc_free( nflags );
	rs.sub[si].nmesh = sub->nummesh;
	for( int j = 0; j < sub->nummesh; j++ )
	{
		const mstudiotexture_t *tx = &ptexture[pskinref[pmesh[j].skinref]];
		const float fs = 1.0f / (float)tx->width, ft = 1.0f / (float)tx->height;
		const short *ptricmds = (const short *)( base + pmesh[j].triindex );
		rc_mcorner_t *out = rc_malloc( sizeof( rc_mcorner_t ) * 3 * Q_max( 1, pmesh[j].numtris + 8 ));
		int ntris = 0, cap = pmesh[j].numtris + 8, i;
		/* the texels come out of TriTexCoord2f with the mesh's skin bound */
		const int keep_w = rc.tri.skinwidth, keep_h = rc.tri.skinheight;
		rc.tri.skinwidth = tx->width; rc.tri.skinheight = tx->height;
		while(( i = *( ptricmds++ )))
		{
			const int fan = i < 0;
			if( fan ) i = -i;
			rc_mcorner_t tv[3];
			int vertcount = 0, n = 0;
			for( ; i > 0; i--, ptricmds += 4 )
			{
				rc_mcorner_t cv;
				cv.vert = (uint16_t)ptricmds[0]; cv.norm = (uint16_t)ptricmds[1];
				if( FBitSet( tx->flags, STUDIO_NF_UV_COORDS )) TriTexCoord2f( HalfToFloat( ptricmds[2] ), HalfToFloat( ptricmds[3] ));
				else TriTexCoord2f( ptricmds[2] * fs, ptricmds[3] * ft );
				cv.s = rc.tri.s; cv.t = rc.tri.t;
				if( ntris + 1 > cap ) continue;
				if( fan )
				{
					tv[vertcount++] = cv;
					if( vertcount >= 3 )
					{
						out[3 * ntris] = tv[0]; out[3 * ntris + 1] = tv[1]; out[3 * ntris + 2] = tv[2]; ntris++;
						tv[1] = tv[2];
						vertcount = 2;
					}
				}
				else
				{
					tv[n] = cv;
					n++; vertcount++;
					if( n == 3 ) n = 0;
					if( vertcount >= 3 )
					{
						if( vertcount & 1 ) { out[3 * ntris] = tv[0]; out[3 * ntris + 1] = tv[1]; out[3 * ntris + 2] = tv[2]; }
						else { out[3 * ntris] = tv[2]; out[3 * ntris + 1] = tv[1]; out[3 * ntris + 2] = tv[0]; }
						ntris++;
					}
				}
			}
		}
		rc.tri.skinwidth = keep_w; rc.tri.skinheight = keep_h;
		rs.sub[si].mesh_first[j] = rs_put( out, sizeof( rc_mcorner_t ) * 3 * Q_max( 1, ntris ), 8 ) / (int32_t)sizeof( rc_mcorner_t );
		rs.sub[si].mesh_ntris[j] = ntris;
		rs.sub[si].mesh_flags[j] = tx->flags;
		rs.sub[si].mesh_tex[j] = tx->index;
		rc_free( out );
	}
	return si;
}

/* slot R_StudioLerpMovement (r_studio.c:491): where a MOVETYPE_STEP entity is between its last two updates */
static void R_StudioLerpMovement( cl_entity_t *e, double time, vec3_t origin, vec3_t angles )
{
	float f = 1.0f;
	if(( time < e->curstate.animtime + 1.0f ) && ( e->curstate.animtime != e->latched.prevanimtime ))
		f = ( time - e->curstate.animtime ) / ( e->curstate.animtime - e->latched.prevanimtime );
	VectorLerp( e->latched.prevorigin, f, e->curstate.origin, origin );
	if( !VectorCompareEpsilon( e->curstate.angles, e->latched.prevangles, ON_EPSILON ))
	{
		vec4_t q, q1, q2;
		AngleQuaternion( e->curstate.angles, q1, false );
		AngleQuaternion( e->latched.prevangles, q2, false );
		QuaternionSlerp( q2, q1, f, q );
		QuaternionAngle( q, angles );
	}
	else VectorCopy( e->curstate.angles, angles );
}

/* R_StudioEstimateFrame (r_studio.c:557) */
static float rc_studio_frame( const cl_entity_t *e, const mstudioseqdesc_t *pseqdesc, double time )
{
	double dfdt = time < e->curstate.animtime ? 0.0 : ( time - e->curstate.animtime ) * e->curstate.framerate * pseqdesc->fps;
	double f = pseqdesc->numframes <= 1 ? 0.0 : ( e->curstate.frame * ( pseqdesc->numframes - 1 )) / 256.0;
	f += dfdt;
	if( pseqdesc->flags & STUDIO_LOOPING )
	{
		if( pseqdesc->numframes > 1 )
			f -= (int)( f / ( pseqdesc->numframes - 1 )) * ( pseqdesc->numframes - 1 );
		if( f < 0 ) f += ( pseqdesc->numframes - 1 );
	}
	else
	{
		if( f >= pseqdesc->numframes - 1.001 ) f = pseqdesc->numframes - 1.001;
		if( f < 0.0 ) f = 0.0;
	}
	return f;
}

static float R_StudioEstimateFrame( cl_entity_t *e, mstudioseqdesc_t *pseqdesc, double time ) { return rc_studio_frame( e, pseqdesc, time ); }

static float rc_studio_interpolant( const cl_entity_t *e )
{
	float dadt = 1.0f;
	if( e->curstate.animtime >= e->latched.prevanimtime + 0.01f )
	{
		dadt = ( gp_cl->time - e->curstate.animtime ) / 0.1f;
		if( dadt > 2.0f ) dadt = 2.0f;
	}
	return dadt;
}

/* R_StudioCalcRotations (r_studio.c:733) with R_StudioCalcBoneAdj (:668) */
static void rc_studio_rotations( const cl_entity_t *e, studiohdr_t *hdr, float pos[][3], vec4_t *q, const mstudioseqdesc_t *pseqdesc, const mstudioanim_t *panim, float f )
{
	if( f > pseqdesc->numframes - 1 ) f = 0.0f;
	else if( f < -0.01f ) f = -0.01f;
	const int frame = (int)f;
	const float dadt = rc_studio_interpolant( e );
	const float s = ( f - frame );
	float adj[MAXSTUDIOCONTROLLERS];
	const mstudiobonecontroller_t *bc = (const mstudiobonecontroller_t *)((const byte *)hdr + hdr->bonecontrollerindex );
	float value = 0.0f;
	for( int j = 0; j < hdr->numbonecontrollers && j < MAXSTUDIOCONTROLLERS; j++ )
	{
		const int i = bc[j].index;
		const byte *c1 = e->curstate.controller, *c2 = e->latched.prevcontroller;
		if( i == STUDIO_MOUTH )
		{
			value = (float)e->mouth.mouthopen / 64.0f;
			value = bound( 0.0f, value, 1.0f );
			value = ( 1.0f - value ) * bc[j].start + value * bc[j].end;
		}
		else if( i < 4 )
		{
			if( FBitSet( bc[j].type, STUDIO_RLOOP ))
			{
				if( abs( c1[i] - c2[i] ) > 128 )
				{
					const int a = ( c1[i] + 128 ) % 256, b = ( c2[i] + 128 ) % 256;
					value = (( a * dadt ) + ( b * ( 1.0f - dadt )) - 128 ) * ( 360.0f / 256.0f ) + bc[j].start;
				}
				else value = (( c1[i] * dadt + ( c2[i] ) * ( 1.0f - dadt ))) * ( 360.0f / 256.0f ) + bc[j].start;
			}
			else
			{
				value = ( c1[i] * dadt + c2[i] * ( 1.0f - dadt )) / 255.0f;
				value = bound( 0.0f, value, 1.0f );
				value = ( 1.0f - value ) * bc[j].start + value * bc[j].end;
			}
		}
		switch( bc[j].type & STUDIO_TYPES )
		{
		case STUDIO_XR: case STUDIO_YR: case STUDIO_ZR: adj[j] = DEG2RAD( value ); break;
		case STUDIO_X: case STUDIO_Y: case STUDIO_Z: adj[j] = value; break;
		}
	}
	const mstudiobone_t *pbone = (const mstudiobone_t *)((const byte *)hdr + hdr->boneindex );
	for( int i = 0; i < hdr->numbones; i++, pbone++, panim++ )
		R_StudioCalcBones( frame, s, pbone, panim, adj, pos[i], q[i] );
	if( pseqdesc->motiontype & STUDIO_X ) pos[pseqdesc->motionbone][0] = 0.0f;
	if( pseqdesc->motiontype & STUDIO_Y ) pos[pseqdesc->motionbone][1] = 0.0f;
	if( pseqdesc->motiontype & STUDIO_Z ) pos[pseqdesc->motionbone][2] = 0.0f;
}

/* R_StudioSetUpTransform (r_studio.c:523): the entity's placement */
static void rc_studio_transform( cl_entity_t *e, matrix3x4 rotation )
{
	vec3_t origin, angles;
	VectorCopy( e->origin, origin ); VectorCopy( e->angles, angles );
	if( e->curstate.movetype == MOVETYPE_STEP && !FBitSet( gp_host->features, ENGINE_COMPUTE_STUDIO_LERP ))
		R_StudioLerpMovement( e, gp_cl->time, origin, angles );
	if( !FBitSet( gp_host->features, ENGINE_COMPENSATE_QUAKE_BUG ))
		angles[PITCH] = -angles[PITCH];
	Matrix3x4_CreateFromEntity( rotation, angles, origin, 1.0f );
}

/* R_StudioFxTransform (r_studio.c:617) as far as it is built: kRenderFxExplode stretches the root bones */
static void rc_studio_fx_transform( const cl_entity_t *e, matrix3x4 bone )
{
	if( e->curstate.renderfx == kRenderFxExplode )
	{
		float scale = 1.0f + ( gp_cl->time - e->curstate.animtime ) * 10.0f;
		if( scale > 2.0f ) scale = 2.0f;
		bone[0][1] *= scale; bone[1][1] *= scale; bone[2][1] *= scale;
	}
}

/* R_StudioSetUpTransform + R_StudioSetupBones (r_studio.c:523,840): bone and light matrices of the entity */
static qboolean rc_studio_skeleton( cl_entity_t *e, model_t *mod, studiohdr_t *hdr, matrix3x4 rotation, matrix3x4 *bones, matrix3x4 *light )
{
	static vec3_t pos[MAXSTUDIOBONES], pos2[MAXSTUDIOBONES], pos3[MAXSTUDIOBONES], pos4[MAXSTUDIOBONES], pos1b[MAXSTUDIOBONES];
	static vec4_t q[MAXSTUDIOBONES], q2[MAXSTUDIOBONES], q3[MAXSTUDIOBONES], q4[MAXSTUDIOBONES], q1b[MAXSTUDIOBONES];
	rc_studio_transform( e, rotation );

	if( e->curstate.sequence >= hdr->numseq ) e->curstate.sequence = 0;
	const int nb = hdr->numbones;
	const double time = gp_cl->time;
	mstudioseqdesc_t *pseqdesc = (mstudioseqdesc_t *)((byte *)hdr + hdr->seqindex ) + e->curstate.sequence;
	const float f = rc_studio_frame( e, pseqdesc, time );
	mstudioanim_t *panim = gEngfuncs.R_StudioGetAnim( hdr, mod, pseqdesc );
	if( !panim )
		return false;
	rc_studio_rotations( e, hdr, pos, q, pseqdesc, panim, f );
	if( pseqdesc->numblends > 1 )
	{
		panim += nb;
		rc_studio_rotations( e, hdr, pos2, q2, pseqdesc, panim, f );
		const float dadt = rc_studio_interpolant( e );
		float s = ( e->curstate.blending[0] * dadt + e->latched.prevblending[0] * ( 1.0f - dadt )) / 255.0f;
		R_StudioSlerpBones( nb, q, pos, q2, pos2, s );
		if( pseqdesc->numblends == 4 )
		{
			panim += nb; rc_studio_rotations( e, hdr, pos3, q3, pseqdesc, panim, f );
			panim += nb; rc_studio_rotations( e, hdr, pos4, q4, pseqdesc, panim, f );
			s = ( e->curstate.blending[0] * dadt + e->latched.prevblending[0] * ( 1.0f - dadt )) / 255.0f;
			R_StudioSlerpBones( nb, q3, pos3, q4, pos4, s );
			s = ( e->curstate.blending[1] * dadt + e->latched.prevblending[1] * ( 1.0f - dadt )) / 255.0f;
			R_StudioSlerpBones( nb, q, pos, q3, pos3, s );
		}
	}
	if( e->latched.sequencetime && ( e->latched.sequencetime + 0.2f > time ) && ( e->latched.prevsequence < hdr->numseq ))
	{
		/* blend from the previous sequence */
		pseqdesc = (mstudioseqdesc_t *)((byte *)hdr + hdr->seqindex ) + e->latched.prevsequence;
		panim = gEngfuncs.R_StudioGetAnim( hdr, mod, pseqdesc );
		if( !panim ) return false;
		rc_studio_rotations( e, hdr, pos1b, q1b, pseqdesc, panim, e->latched.prevframe );
		if( pseqdesc->numblends > 1 )
		{
			panim += nb; rc_studio_rotations( e, hdr, pos2, q2, pseqdesc, panim, e->latched.prevframe );
			float s = ( e->latched.prevseqblending[0] ) / 255.0f;
			R_StudioSlerpBones( nb, q1b, pos1b, q2, pos2, s );
			if( pseqdesc->numblends == 4 )
			{
				panim += nb; rc_studio_rotations( e, hdr, pos3, q3, pseqdesc, panim, e->latched.prevframe );
				panim += nb; rc_studio_rotations( e, hdr, pos4, q4, pseqdesc, panim, e->latched.prevframe );
				s = ( e->latched.prevseqblending[0] ) / 255.0f;
				R_StudioSlerpBones( nb, q3, pos3, q4, pos4, s );
				s = ( e->latched.prevseqblending[1] ) / 255.0f;
				R_StudioSlerpBones( nb, q1b, pos1b, q3, pos3, s );
			}
		}
		const float s = 1.0f - ( time - e->latched.sequencetime ) / 0.2f;
		R_StudioSlerpBones( nb, q, pos, q1b, pos1b, s );
	}
	else e->latched.prevframe = f;

	const mstudiobone_t *pbones = (const mstudiobone_t *)((const byte *)hdr + hdr->boneindex );
	for( int i = 0; i < nb; i++ )
	{
		matrix3x4 bonematrix;
		Matrix3x4_FromOriginQuat( bonematrix, q[i], pos[i] );
		if( pbones[i].parent == -1 )
		{
			Matrix3x4_ConcatTransforms( bones[i], rotation, bonematrix );
			Matrix3x4_Copy( light[i], bones[i] );
			rc_studio_fx_transform( e, bones[i] );
		}
		else
		{
			Matrix3x4_ConcatTransforms( bones[i], bones[pbones[i].parent], bonematrix );
			Matrix3x4_ConcatTransforms( light[i], light[pbones[i].parent], bonematrix );
		}
	}
	return true;
}

/*
 * A model that rides on another one (MOVETYPE_FOLLOW: a carried weapon, a rider) - R_DrawStudioModel's branch
 * (r_studio.c:2869) + R_StudioMergeBones (:780): a skeleton is set up for the parent entity and cached, the follower takes
 * the parent's place, and every bone of it that carries the name of a cached bone takes that bone's matrices; the others
 * hang off them with the follower's own animation.
 */
static qboolean rc_studio_follow_skeleton( cl_entity_t *e, studiohdr_t *hdr, matrix3x4 rotation, matrix3x4 *bones, matrix3x4 *light )
{
	static matrix3x4 pbone[MAXSTUDIOBONES], plight[MAXSTUDIOBONES];
	static vec3_t pos[MAXSTUDIOBONES];
	static vec4_t q[MAXSTUDIOBONES];
	cl_entity_t *parent = ( rc.entities && e->curstate.aiment > 0 && (uint)e->curstate.aiment < rc.max_entities ) ? &rc.entities[e->curstate.aiment] : NULL;
	if( !parent || !parent->model || parent->model->type != mod_studio )
		return false;
	/* R_FindParentEntity: the parent has to be in this frame's lists */
	cl_entity_t *found = NULL;
	for( int i = 0; i < rc.num_solid && !found; i++ ) if( rc.solid_ents[i]->index == e->curstate.aiment ) found = rc.solid_ents[i];
	for( int i = 0; i < rc.num_trans && !found; i++ ) if( rc.trans_ents[i]->index == e->curstate.aiment ) found = rc.trans_ents[i];
	if( !found || found->player || found->curstate.movetype == MOVETYPE_FOLLOW || found->curstate.renderfx == kRenderFxDeadPlayer )
		return false; /* (players and chains of followers are not built) */
	/* the parent's pass runs with RI.currentmodel still the follower's (R_DrawEntitiesOnList set it, r_main.c:548, and
	 * R_DrawStudioModel changes RI.currententity only): the skeleton that is cached and merged from is the FOLLOWER's
	 * model posed by the parent's entity state - origin, angles, sequence, frame, controllers */
	studiohdr_t *phdr = hdr;
	matrix3x4 prot;
	if( !rc_studio_skeleton( found, e->model, phdr, prot, pbone, plight ))
		return false;
	VectorCopy( found->curstate.origin, e->curstate.origin );
	VectorCopy( found->origin, e->origin );

	rc_studio_transform( e, rotation );
	if( e->curstate.sequence >= hdr->numseq ) e->curstate.sequence = 0;
	mstudioseqdesc_t *pseqdesc = (mstudioseqdesc_t *)((byte *)hdr + hdr->seqindex ) + e->curstate.sequence;
	const float f = rc_studio_frame( e, pseqdesc, gp_cl->time );
	mstudioanim_t *panim = gEngfuncs.R_StudioGetAnim( hdr, e->model, pseqdesc );
	if( !panim )
		return false;
	rc_studio_rotations( e, hdr, pos, q, pseqdesc, panim, f );
	const mstudiobone_t *mybones = (const mstudiobone_t *)((const byte *)hdr + hdr->boneindex );
	const mstudiobone_t *parbones = (const mstudiobone_t *)((const byte *)phdr + phdr->boneindex );
	for( int i = 0; i < hdr->numbones; i++ )
	{
		int j;
		for( j = 0; j < phdr->numbones; j++ )
			if( !strcasecmp( mybones[i].name, parbones[j].name ))
			{
				Matrix3x4_Copy( bones[i], pbone[j] );
				Matrix3x4_Copy( light[i], plight[j] );
				break;
			}
		if( j < phdr->numbones )
			continue;
		matrix3x4 bonematrix;
		Matrix3x4_FromOriginQuat( bonematrix, q[i], pos[i] );
		if( mybones[i].parent == -1 )
		{
			Matrix3x4_ConcatTransforms( bones[i], rotation, bonematrix );
			Matrix3x4_Copy( light[i], bones[i] );
			rc_studio_fx_transform( e, bones[i] );
		}
		else
		{
			Matrix3x4_ConcatTransforms( bones[i], bones[mybones[i].parent], bonematrix );
			Matrix3x4_ConcatTransforms( light[i], light[mybones[i].parent], bonematrix );
		}
	}
	return true;
}

/* R_RecursiveLightPoint (ref/common/ref_light.c:222) over the world's nodes: the first lightmapped face a segment hits */
static vec3_t rs_lightspot, rs_lightvec;

static qboolean rc_light_trace( model_t *model, mnode_t *node, const vec3_t start, const vec3_t end, uint *cv )
{
	while( node && node->contents >= 0 )
	{
		const float front = PlaneDiff( start, node->plane ), back = PlaneDiff( end, node->plane );
		const int side = front < 0;
		if(( back < 0 ) == side )
		{
			node = node_child( node, side, model );
			continue;
		}
		const float frac = front / ( front - back );
		vec3_t mid;
		VectorLerp( start, frac, end, mid );
		if( rc_light_trace( model, node_child( node, side, model ), start, mid, cv ))
			return true;
		const int firstsurface = node_firstsurface( node, model ), numsurfaces = node_numsurfaces( node, model );
		VectorCopy( mid, rs_lightspot );
		for( int i = 0; i < numsurfaces; i++ )
		{
			const msurface_t *surf = &model->surfaces[firstsurface + i];
			const mextrasurf_t *info = surf->info;
			if( FBitSet( surf->flags, SURF_DRAWTILED ))
				continue;
			const float s = DotProduct( mid, info->lmvecs[0] ) + info->lmvecs[0][3];
			const float t = DotProduct( mid, info->lmvecs[1] ) + info->lmvecs[1][3];
			if( s < info->lightmapmins[0] || t < info->lightmapmins[1] )
				continue;
			float ds = s - info->lightmapmins[0], dt = t - info->lightmapmins[1];
			if( ds > info->lightextents[0] || dt > info->lightextents[1] )
				continue;
			cv[0] = cv[1] = cv[2] = 0;
			if( !surf->samples )
				return true;
			const int sample_size = gEngfuncs.Mod_SampleSizeForFace( surf );
			const int smax = ( info->lightextents[0] / sample_size ) + 1, tmax = ( info->lightextents[1] / sample_size ) + 1;
			ds /= sample_size; dt /= sample_size;
			const color24 *lm = surf->samples + Q_rint( dt ) * smax + Q_rint( ds );
			if( info->deluxemap && !rc.warned[6] ) { rc.warned[6] = true; gEngfuncs.Con_Reportf( "ref_cuda: deluxe maps are not used for entity light vectors\n" ); }
			for( int map = 0; map < MAXLIGHTMAPS && surf->styles[map] != 255; map++ )
			{
				const uint scale = rc.lightstylevalue[surf->styles[map]];
				cv[0] += lm->r * scale; cv[1] += lm->g * scale; cv[2] += lm->b * scale;
				lm += smax * tmax;
			}
			return true;
		}
		/* go down the back side */
		return rc_light_trace( model, node_child( node, !side, model ), mid, end, cv );
	}
	cv[0] = cv[1] = cv[2] = 0;
	return false;
}

/* R_LightVec (ref_light.c:450) against the world (the first physent): light colour 0..255 per channel */
static void rc_light_vec( const vec3_t start, const vec3_t end, qboolean want_vec, uint *light )
{
	light[0] = light[1] = light[2] = 0;
	VectorClear( rs_lightspot ); VectorClear( rs_lightvec );
	model_t *world = gp_cl->models[1];
	if( !world || !world->lightdata ) { light[0] = light[1] = light[2] = 255; return; }
	for( int pass = 0; pass < 2; pass++ )
	{
		uint cv[3];
		mnode_t *pnodes = &world->nodes[world->hulls[0].firstclipnode];
		vec3_t offset, s_l, e_l;
		VectorSubtract( world->hulls[0].clip_mins, vec3_origin, offset );
		VectorSubtract( pass ? end : start, offset, s_l );
		VectorSubtract( pass ? start : end, offset, e_l );
		if( rc_light_trace( world, pnodes, s_l, e_l, cv ))
		{
			light[0] = Q_min(( cv[0] >> 8 ), 255 ); light[1] = Q_min(( cv[1] >> 8 ), 255 ); light[2] = Q_min(( cv[2] >> 8 ), 255 );
		}
		/* nothing lit below: r_lighting_extended tries the ceiling when the caller asked for the light vector */
		if(( light[0] + light[1] + light[2] ) != 0 || !want_vec || !gEngfuncs.pfnGetCvarFloat( "r_lighting_extended" ))
			break;
	}
}

/* R_EntityDynamicLight (ref_light.c:498): ambient / shade light, colour and direction for a model */
/* (alight_t.ambientlight / shadelight are ints, common/com_model.h:405: the float results are truncated into them) */
static void rc_entity_light( cl_entity_t *ent, int *ambientlight, int *shadelight, vec3_t color, vec3_t plightvec )
{
	if( r_fullbright->value || FBitSet( ent->curstate.effects, EF_FULLBRIGHT ))
	{
		*shadelight = 0; *ambientlight = 192;
		VectorSet( plightvec, 0.0f, 0.0f, -1.0f ); VectorSet( color, 1.0f, 1.0f, 1.0f );
		return;
	}
	const movevars_t *mv = (const movevars_t *)gEngfuncs.EngineGetParm( PARM_GET_MOVEVARS_PTR, 0 );
	vec3_t lightDir, origin, vecSrc, vecEnd;
	if( FBitSet( ent->curstate.effects, EF_INVLIGHT )) VectorSet( lightDir, 0.0f, 0.0f, 1.0f );
	else VectorSet( lightDir, 0.0f, 0.0f, -1.0f );
	VectorCopy( ent->origin, origin );
	VectorSet( vecSrc, origin[0], origin[1], origin[2] - lightDir[2] * 8.0f );
	uint light[3] = { 0, 0, 0 };
	if( mv && ( mv->skycolor[0] + mv->skycolor[1] + mv->skycolor[2] ) != 0 && !rc.warned[7] )
	{
		rc.warned[7] = true;
		gEngfuncs.Con_Reportf( "ref_cuda: sky light on models (movevars skycolor) is not built\n" );
	}
	VectorScale( lightDir, 2048.0f, vecEnd );
	VectorAdd( vecEnd, vecSrc, vecEnd );
	rc_light_vec( vecSrc, vecEnd, true, light );
	/* no deluxe data: the light vector is null and the direction comes from the gradient of the floor light */
	{
		float grad[4];
		uint g[3];
		vecSrc[0] -= 16.0f; vecSrc[1] -= 16.0f; vecEnd[0] -= 16.0f; vecEnd[1] -= 16.0f;
		rc_light_vec( vecSrc, vecEnd, false, g ); grad[0] = ( g[0] + g[1] + g[2] ) / 768.0f;
		vecSrc[0] += 32.0f; vecEnd[0] += 32.0f;
		rc_light_vec( vecSrc, vecEnd, false, g ); grad[1] = ( g[0] + g[1] + g[2] ) / 768.0f;
		vecSrc[1] += 32.0f; vecEnd[1] += 32.0f;
		rc_light_vec( vecSrc, vecEnd, false, g ); grad[2] = ( g[0] + g[1] + g[2] ) / 768.0f;
		vecSrc[0] -= 32.0f; vecEnd[0] -= 32.0f;
		rc_light_vec( vecSrc, vecEnd, false, g ); grad[3] = ( g[0] + g[1] + g[2] ) / 768.0f;
		lightDir[0] = grad[0] - grad[1] - grad[2] + grad[3];
		lightDir[1] = grad[1] + grad[0] - grad[2] - grad[3];
		VectorNormalize( lightDir );
	}
	if( ent->curstate.renderfx == kRenderFxLightMultiplier && ent->curstate.iuser4 != 10 )
	{
		light[0] *= ent->curstate.iuser4 / 10.0f; light[1] *= ent->curstate.iuser4 / 10.0f; light[2] *= ent->curstate.iuser4 / 10.0f;
	}
	vec3_t finalLight = { light[0], light[1], light[2] };
	ent->cvFloorColor.r = light[0]; ent->cvFloorColor.g = light[1]; ent->cvFloorColor.b = light[2]; ent->cvFloorColor.a = 0;
	float total = Q_max( Q_max( light[0], light[1] ), light[2] );
	if( total == 0.0f ) total = 1.0f;
	VectorScale( lightDir, total, lightDir );
	const dlight_t *dls = (const dlight_t *)gEngfuncs.EngineGetParm( PARM_GET_DLIGHTS_PTR, 0 );
	for( uint lnum = 0; dls && lnum < MAX_DLIGHTS; lnum++ )
	{
		const dlight_t *dl = &dls[lnum];
		if( dl->die < gp_cl->time || !r_dynamic->value )
			continue;
		vec3_t dist;
		VectorSubtract( ent->origin, dl->origin, dist );
		const float radius = VectorLength( dist );
		const float add = ( dl->radius - radius );
		if( add > 0.0f )
		{
			total += add;
			if( radius > 1.0f ) VectorScale( dist, ( add / radius ), dist );
			else VectorScale( dist, add, dist );
			VectorAdd( lightDir, dist, lightDir );
			finalLight[0] += dl->color.r * ( add / 256.0f );
			finalLight[1] += dl->color.g * ( add / 256.0f );
			finalLight[2] += dl->color.b * ( add / 256.0f );
		}
	}
	float scale;
	if( ent->model->type == mod_studio && FBitSet( ent->model->flags, STUDIO_AMBIENT_LIGHT )) scale = 0.6f;
	else scale = bound( 0.75f, v_direct->value, 1.0f );
	VectorScale( lightDir, scale, lightDir );
	*shadelight = VectorLength( lightDir );
	*ambientlight = total - *shadelight;
	total = Q_max( Q_max( finalLight[0], finalLight[1] ), finalLight[2] );
	if( total > 0.0f )
	{
		color[0] = finalLight[0] * ( 1.0f / total ); color[1] = finalLight[1] * ( 1.0f / total ); color[2] = finalLight[2] * ( 1.0f / total );
	}
	else VectorSet( color, 1.0f, 1.0f, 1.0f );
	if( *ambientlight > 128 ) *ambientlight = 128;
	if( *ambientlight + *shadelight > 255 ) *shadelight = 255 - *ambientlight;
	VectorNormalize2( lightDir, plightvec );
}

static void rc_add_studio( cl_entity_t *e )
{
	studiohdr_t *hdr = (studiohdr_t *)gEngfuncs.Mod_Extradata( mod_studio, e->model );
	if( !hdr )
		return;
	if( e->player || e->curstate.renderfx == kRenderFxGlowShell || e->curstate.renderfx == kRenderFxDeadPlayer
		|| FBitSet( hdr->flags, STUDIO_HAS_BONEWEIGHTS ) || hdr->numbones > MAXSTUDIOBONES )
	{
		if( !rc.warned[5] ) { rc.warned[5] = true; gEngfuncs.Con_Reportf( "ref_cuda: player models, glow shells and bone-weighted models are not drawn yet\n" ); }
		return;
	}
	static matrix3x4 bones[MAXSTUDIOBONES], light[MAXSTUDIOBONES];
	matrix3x4 rotation;
	if( e->curstate.movetype == MOVETYPE_FOLLOW )
	{
		if( !rc_studio_follow_skeleton( e, hdr, rotation, bones, light ))
			return;
		if( hdr->numbodyparts == 0 )
			return;
	}
	else
	{
		if( hdr->numbodyparts == 0 )
			return;
		if( !rc_studio_skeleton( e, e->model, hdr, rotation, bones, light ))
			return;
	}
	/* R_StudioCalcAttachments (r_studio.c:1212): where the model's attachment points are this frame - sprites and models
	 * that follow the entity (MOVETYPE_FOLLOW, body = attachment) are placed there */
	{
		const mstudioattachment_t *att = (const mstudioattachment_t *)((const byte *)hdr + hdr->attachmentindex );
		for( int i = 0; i < Q_min( MAXSTUDIOATTACHMENTS, hdr->numattachments ); i++ )
			Matrix3x4_VectorTransform( light[att[i].bone], att[i].org, e->attachment[i] );
	}
	/* attachments of a model that has none: the entity origin (R_StudioClientEvents, r_studio.c:2172) */
	if( hdr->numattachments <= 0 && gp_cl->time != gp_cl->oldtime )
		for( int i = 0; i < 4; i++ ) Matrix3x4_OriginFromMatrix( rotation, e->attachment[i] );
	int ambientlight, shadelight;
	vec3_t lightcolor, plightvec;
	rc_entity_light( e, &ambientlight, &shadelight, lightcolor, plightvec );

	/* the entity's bones and bone-space light vectors (R_StudioSetupLighting, r_studio.c:1361) join the batch */
	const int nb = hdr->numbones, bone0 = rs.nbones;
	rs.bones = rc_grow( rs.bones, &rs.maxbones, ( bone0 + nb ) * 12, sizeof( float ));
	rs.blight = rc_grow( rs.blight, &rs.maxblight, ( bone0 + nb ) * 3, sizeof( float ));
	const float scale = e->curstate.scale;
	for( int i = 0; i < nb; i++ )
	{
		memcpy( rs.bones + ( bone0 + i ) * 12, bones[i], sizeof( matrix3x4 ));
		vec3_t bl;
		Matrix3x4_VectorIRotate( light[i], plightvec, bl );
		if( scale > 1.0f ) VectorNormalize( bl );
		VectorCopy( bl, rs.blight + ( bone0 + i ) * 3 );
	}
	rs.nbones += nb;

	int rendermode = e->curstate.rendermode;
	if( rendermode > kRenderTransAdd ) rendermode = 0; /* R_StudioSetupRenderer's clamp applies to g_studio.rendermode only; the bind uses the entity's */
	rendermode = e->curstate.rendermode;
	const mstudiobodyparts_t *bp0 = (const mstudiobodyparts_t *)((const byte *)hdr + hdr->bodypartindex );
	for( int b = 0; b < hdr->numbodyparts; b++ )
	{
		const mstudiobodyparts_t *bp = bp0 + b;
		int index = e->curstate.body / bp->base;
		index = index % bp->nummodels;
		mstudiomodel_t *sub = (mstudiomodel_t *)((byte *)hdr + bp->modelindex ) + index;
		const int si = rs_submodel( hdr, sub, e->curstate.skin );
		if( si < 0 )
			continue;
		rs.parts = rc_grow( rs.parts, &rs.maxparts, rs.nparts + 1, sizeof( rc_spart_t ));
		rc_spart_t *pt = &rs.parts[rs.nparts++];
		memset( pt, 0, sizeof( *pt ));
		pt->vert_base = rs.totalverts; pt->numverts = sub->numverts;
		pt->light_base = rs.totalnorms; pt->numnorms = sub->numnorms;
		pt->sverts = rs.sub[si].sverts; pt->svertbone = rs.sub[si].svertbone;
		pt->snorms = rs.sub[si].snorms; pt->snormbone = rs.sub[si].snormbone; pt->snormflags = rs.sub[si].snormflags;
		pt->bones = bone0; pt->blightvec = bone0;
		pt->ambientlight = ambientlight; pt->shadelight = shadelight; VectorCopy( lightcolor, pt->lightcolor );
		pt->blend = rc.tri.blend; pt->transadd = rendermode == kRenderTransAdd;
		rs.totalverts += sub->numverts; rs.totalnorms += sub->numnorms;

		/* GL_StudioSetRenderMode + R_StudioDrawPoints' mesh loop: one draw item per mesh, skin bound under the entity's mode */
		GL_SetRenderMode( rendermode );
		int order[MAXSTUDIOMESHES];
		for( int j = 0; j < rs.sub[si].nmesh; j++ ) order[j] = j;
		/* r_studio_sort_textures: opaque meshes first, then masked, then additive (R_StudioMeshCompare, r_studio.c:1667);
		 * insertion sort keeps equal ones in file order, as the reference's comparison of unequal flags needs */
		for( int a = 1; a < rs.sub[si].nmesh; a++ )
			for( int c = a; c > 0; c-- )
			{
				const int fa = rs.sub[si].mesh_flags[order[c - 1]], fb = rs.sub[si].mesh_flags[order[c]];
				const int ka = FBitSet( fa, STUDIO_NF_ADDITIVE ) ? 2 : FBitSet( fa, STUDIO_NF_MASKED ) ? 1 : 0;
				const int kb = FBitSet( fb, STUDIO_NF_ADDITIVE ) ? 2 : FBitSet( fb, STUDIO_NF_MASKED ) ? 1 : 0;
				if( ka <= kb ) break;
				const int tmp = order[c]; order[c] = order[c - 1]; order[c - 1] = tmp;
			}
		for( int jj = 0; jj < rs.sub[si].nmesh; jj++ )
		{
			const int j = order[jj];
			GL_Bind( XASH_TEXTURE0, rs.sub[si].mesh_tex[j] );
			rc_item_t *it = rc_new_item( RC_ITEM_MESH );
			it->first = rs.sub[si].mesh_first[j];
			it->count = rs.sub[si].mesh_ntris[j];
			it->vert_base = pt->vert_base; it->light_base = pt->light_base;
			it->part = (int)( pt - rs.parts );
			it->chrome = FBitSet( rs.sub[si].mesh_flags[j], STUDIO_NF_CHROME ) ? 1 : 0;
		}
	}
}

/* ---------------------------------------------------------------------------------------------- */
/* opaque brush entities: the per-entity, view-independent part of R_DrawBEntitiesOnList (ref/soft/r_main.c:781) - the
 * rotated bounding box (RotatedBBox, :723) and the rotation matrix R_RotateBmodel builds (r_bsp.c:97-160, sin / cos on
 * the host) - for the device, which clips every face to the view and to the world BSP per view (rc_core.h A13)           */

static void rc_concat_rotations( float in1[3][3], float in2[3][3], float out[3][3] )
{
	for( int i = 0; i < 3; i++ )
		for( int j = 0; j < 3; j++ )
			out[i][j] = in1[i][0] * in2[0][j] + in1[i][1] * in2[1][j] + in1[i][2] * in2[2][j];
}

static rc_bent_t rc_bents[RC_MAX_BENTS];
static int rc_nbents;

/* -> index of the new record, -1 = not a drawable brush model / table full */
static int rc_make_bent( cl_entity_t *e )
{
	model_t *mod = e->model;
	if( !mod || mod->nummodelsurfaces == 0 || mod->type != mod_brush || rc_nbents >= RC_MAX_BENTS )
		return -1;
	rc_bent_t *b = &rc_bents[rc_nbents];
	memset( b, 0, sizeof( *b ));
	/* RotatedBBox */
	vec3_t tmins, tmaxs;
	if( !e->angles[0] && !e->angles[1] && !e->angles[2] )
	{
		VectorCopy( mod->mins, tmins ); VectorCopy( mod->maxs, tmaxs );
	}
	else
	{
		vec3_t forward, right, up, tmp, v;
		for( int k = 0; k < 3; k++ ) { tmins[k] = 99999; tmaxs[k] = -99999; }
		AngleVectors( e->angles, forward, right, up );
		for( int c = 0; c < 8; c++ )
		{
			tmp[0] = ( c & 1 ) ? mod->mins[0] : mod->maxs[0];
			tmp[1] = ( c & 2 ) ? mod->mins[1] : mod->maxs[1];
			tmp[2] = ( c & 4 ) ? mod->mins[2] : mod->maxs[2];
			VectorScale( forward, tmp[0], v );
			VectorMA( v, -tmp[1], right, v );
			VectorMA( v, tmp[2], up, v );
			for( int j = 0; j < 3; j++ )
			{
				if( v[j] < tmins[j] ) tmins[j] = v[j];
				if( v[j] > tmaxs[j] ) tmaxs[j] = v[j];
			}
		}
	}
	VectorAdd( tmins, e->origin, b->minmaxs );
	VectorAdd( tmaxs, e->origin, ( b->minmaxs + 3 ));
	VectorCopy( e->origin, b->origin );
	/* R_RotateBmodel: yaw, pitch, roll */
	float angle, sn, cs, temp1[3][3], temp2[3][3], temp3[3][3];
	angle = e->angles[YAW]; angle = angle * M_PI_F * 2 / 360.0f; sn = sin( angle ); cs = cos( angle );
	temp1[0][0] = cs; temp1[0][1] = sn; temp1[0][2] = 0; temp1[1][0] = -sn; temp1[1][1] = cs; temp1[1][2] = 0; temp1[2][0] = 0; temp1[2][1] = 0; temp1[2][2] = 1;
	angle = e->angles[PITCH]; angle = angle * M_PI_F * 2 / 360.0f; sn = sin( angle ); cs = cos( angle );
	temp2[0][0] = cs; temp2[0][1] = 0; temp2[0][2] = -sn; temp2[1][0] = 0; temp2[1][1] = 1; temp2[1][2] = 0; temp2[2][0] = sn; temp2[2][1] = 0; temp2[2][2] = cs;
	rc_concat_rotations( temp2, temp1, temp3 );
	angle = e->angles[ROLL]; angle = angle * M_PI_F * 2 / 360.0f; sn = sin( angle ); cs = cos( angle );
	temp1[0][0] = 1; temp1[0][1] = 0; temp1[0][2] = 0; temp1[1][0] = 0; temp1[1][1] = cs; temp1[1][2] = sn; temp1[2][0] = 0; temp1[2][1] = -sn; temp1[2][2] = cs;
	rc_concat_rotations( temp1, temp3, b->rot );
	b->firstface = mod->firstmodelsurface; b->numfaces = mod->nummodelsurfaces;
	return rc_nbents++;
}

/* the frame's edge_entities (opaque); the translucent ones join the table as rc_build_items reaches them */
static void rc_build_bents( void )
{
	rc_nbents = 0;
	for( int i = 0; i < rc.num_edge; i++ )
		rc_make_bent( rc.edge_ents[i] );
}

/* R_DrawBrushModel (r_main.c:861) for a translucent entity: a draw item in the sorted sequence; the span drawer and the
 * alpha are D_AlphaSurf's (r_edge.c:931-960) */
static void rc_add_tbrush( cl_entity_t *e )
{
	const int bi = rc_make_bent( e );
	if( bi < 0 )
		return;
	rc_bent_t *b = &rc_bents[bi];
	int alpha = e->curstate.renderamt * 7 / 255;
	if( alpha <= 0 && e->curstate.renderamt > 0 )
		alpha = 1;
	b->alpha = alpha;
	b->trans = e->curstate.rendermode == kRenderTransAlpha ? RC_TB_ALPHATEST : e->curstate.rendermode == kRenderTransAdd ? RC_TB_ADD : RC_TB_BLEND;
	rc_item_t *it = rc_new_item( RC_ITEM_BRUSH );
	it->first = bi; it->count = 0;
	it->tex_ofs = 0; /* no skin: the pieces sample the surface cache */
	it->alpha = alpha;
}

/* LightToTexGamma (ref/soft/r_local.h:518-523) reads the engine's live table, or nothing in linear gamma space */
static void rc_resolve_lightgamma( void )
{
	static uint16_t identity[1024];
	rc.lightgamma = (const uint16_t *)gEngfuncs.EngineGetParm( PARM_GET_LIGHTGAMMATABLE_PTR, 0 );
	if( !rc.lightgamma || FBitSet( gp_host->features, ENGINE_LINEAR_GAMMA_SPACE ))
	{
		for( int i = 0; i < 1024; i++ ) identity[i] = i;
		rc.lightgamma = identity;
	}
}

/* the device keeps a copy of the 2 KB table: V_CheckGamma rebuilds the engine's in place (gamma.c:135-139) and
 * host.features can change when a game loads - resend it when either happened, and relight every cache block */
static void rc_sync_lightgamma( void )
{
	rc_resolve_lightgamma();
	if( !memcmp( rc.lightgamma_sent, rc.lightgamma, sizeof( rc.lightgamma_sent )))
		return;
	memcpy( rc.lightgamma_sent, rc.lightgamma, sizeof( rc.lightgamma_sent ));
	RC_CHECK_DEV( rc_dev_upload_luts( rc.colormap, rc.lightgamma, rc.sintable ));
	RC_CHECK_DEV( rc_dev_flush_cache());
}

static void rc_send_decals( void );
static void rc_prepare_batch( void )
{
	if( !rc.scene_valid )
		gEngfuncs.Host_Error( "ref_cuda: render without a world (R_NewMap not called)\n" );
	rc_sync_lightgamma();
	rc_sync_textures();
	rc_resolve_texnums( false );
	rc_send_decals();
	RC_CHECK_DEV( rc_dev_set_lightstyles( rc.lightstylevalue ));
	rc_push_dlights();
	rc_build_bents();
	rc_build_items();
	RC_CHECK_DEV( rc_dev_set_bents( rc_bents, rc_nbents ));
}

/*
 * R_SpriteOccluded + R_SpriteGlowBlend (ref/soft/r_sprite.c:59-117) for every (view, glow sprite) of the batch: the scissor
 * test on R_WorldToScreen's answer (clip-space coordinates against the viewport rectangle in pixels, as the reference
 * writes it), the engine's sight line from the eye (EV_VisTraceLine), brightness and size from the distance, and the
 * colour _TriColor4f derives from tr.blend * brightness.
 */
static void rc_build_glow( void )
{
	if( !rc.nglow || !rc.nviews )
	{
		RC_CHECK_DEV( rc_dev_set_glow( NULL, 0, 0 ));
		return;
	}
	rc.glowrecs = rc_grow( rc.glowrecs, &rc.maxglowrecs, rc.nglow * rc.nviews, sizeof( rc_glow_t ));
	const int keepmode = rc.tri.rendermode;
	rc.tri.rendermode = kRenderGlow;
	for( int v = 0; v < rc.nviews; v++ )
	{
		const rc_view_t *vw = &rc.views[v];
		for( int k = 0; k < rc.nglow; k++ )
		{
			const struct rc_glowsrc_s *g = &rc.glowsrc[k];
			rc_glow_t *out = &rc.glowrecs[(size_t)v * rc.nglow + k];
			memset( out, 0, sizeof( *out ));
			vec3_t org, eye, scr;
			VectorCopy( g->origin, org ); VectorCopy( vw->origin, eye );
			rc_world_to_screen_m( vw->wvp, org, scr );
			if( scr[0] < vw->vrect[0] || scr[0] > vw->vrect[0] + vw->vrect[2] )
				continue; /* do scissor */
			if( scr[1] < vw->vrect[1] || scr[1] > vw->vrect[1] + vw->vrect[3] )
				continue;
			float scale = g->scale, blend;
			vec3_t glowDist;
			VectorSubtract( org, eye, glowDist );
			const float dist = VectorLength( glowDist );
			pmtrace_t *tr = gEngfuncs.EV_VisTraceLine( eye, org, r_traceglow.value ? PM_GLASS_IGNORE : ( PM_GLASS_IGNORE | PM_STUDIO_IGNORE ));
			if(( 1.0f - tr->fraction ) * dist > 8.0f )
				blend = 0.0f;
			else if( g->renderfx == kRenderFxNoDissipation )
				blend = 1.0f;
			else
			{
				float brightness = 19000.0f / ( dist * dist ); /* GLARE_FALLOFF */
				blend = bound( 0.05f, brightness, 1.0f );
				scale *= dist * ( 1.0f / 200.0f );
			}
			if( blend <= 0.01f )
				continue; /* faded */
			rc_tri_color4f( g->color[0], g->color[1], g->color[2], g->blend * blend );
			out->visible = 1; out->scale = scale;
			out->light = (int32_t)( rc.tri.light << 8 ); out->color = rc.tri.color; out->alpha = rc.tri.alpha;
		}
	}
	rc.tri.rendermode = keepmode;
	RC_CHECK_DEV( rc_dev_set_glow( rc.glowrecs, rc.nglow, rc.nviews ));
}

static void rc_queue_view( const ref_viewpass_t *rvp )
{
	if( rc.nviews == rc.maxviews )
	{
		int n = rc.maxviews ? rc.maxviews * 2 : 64;
		rc.views = gEngfuncs._Mem_Realloc( rc.pool, rc.views, sizeof( rc_view_t ) * n, true, __FILE__, __LINE__ );
		rc.maxviews = n;
	}
	rc_setup_view( rvp, &rc.views[rc.nviews++] );
}

static void R_Set2DMode( qboolean enable );
static void rc_flush_2d( int view );
static void R_BeginFrame( qboolean clearScene )
{
	(void)clearScene;
	rc.nviews = 0; rc.ncmds2d = 0;
	rc.batch_wvp = true; /* the frame's views are set up before its entities are known */
	R_Set2DMode( true );
}

/* GL_RenderFrame slot == R_RenderFrame (ref/soft/r_main.c:1141): here the view is only recorded */
static void R_RenderFrame( const ref_viewpass_t *rvp )
{
	if( !FBitSet( rvp->flags, RF_DRAW_WORLD ))
		return; /* menu / model-viewer passes have no world; entity-only views arrive with the studio path */
	rc_queue_view( rvp );
}

static void R_RenderScene( void ) {}

/* pixel_t -> display format of the engine's software surface.  The table is R_BuildScreenMap's
 * (ref/soft/r_glblit.c:347-407): entry (i << 8 | j) = major(i) | minor(j), every 5/6/5-bit channel value c
 * rescaled as c * 2^bits / (31 | 63 | 31) and shifted to its mask.  The per-pixel lookup (R_BlitScreen,
 * r_glblit.c:591-721) is k_blit on the device. */
static uint rc_mask_shift( uint m ) { uint s = 0; if( !m ) return 0; while( !( m & 1 )) { m >>= 1; s++; } return s; }
static uint rc_mask_bits( uint m ) { uint c = 0; while( m ) { c += m & 1; m >>= 1; } return c; }
#define RC_BIT_TO( v, from, to ) (((( v ) >> ( from )) & 1u ) << ( to ))

static void rc_build_screenmap( uint32_t *table, uint rmask, uint gmask, uint bmask )
{
	const uint rs = rc_mask_shift( rmask ), gs = rc_mask_shift( gmask ), bs = rc_mask_shift( bmask );
	const uint rm = 1u << rc_mask_bits( rmask ), gm = 1u << rc_mask_bits( gmask ), bm = 1u << rc_mask_bits( bmask );
	uint32_t major[256], minor[256];
	for( uint i = 0; i < 256; i++ )
	{
		/* major byte: 3-3-2 bits of r, g, b at the top of the 5/6/5-bit channels */
		uint r = (( i >> 5 ) << 2 ) & 31u, g = (( i >> 2 ) << 3 ) & 63u, b = ( i << 3 ) & 31u;
		major[i] = ( r * rm / 31u ) << rs | ( g * gm / 63u ) << gs | ( b * bm / 31u ) << bs;
		/* minor byte: the low channel bits, interleaved G B R G B R G B from bit 7 down */
		r = RC_BIT_TO( i, 5, 1 ) | RC_BIT_TO( i, 2, 0 );
		g = RC_BIT_TO( i, 7, 2 ) | RC_BIT_TO( i, 4, 1 ) | RC_BIT_TO( i, 1, 0 );
		b = RC_BIT_TO( i, 6, 2 ) | RC_BIT_TO( i, 3, 1 ) | RC_BIT_TO( i, 0, 0 );
		minor[i] = ( r * rm / 31u ) << rs | ( g * gm / 63u ) << gs | ( b * bm / 31u ) << bs;
	}
	for( uint i = 0; i < 256; i++ )
		for( uint j = 0; j < 256; j++ )
			table[( i << 8 ) | j] = major[i] | minor[j];
}

static int rc_set_display_format( uint bpp, uint rmask, uint gmask, uint bmask )
{
	if( rc.map_valid && rc.map_bpp == bpp && rc.map_r == rmask && rc.map_g == gmask && rc.map_b == bmask )
		return 0;
	uint32_t *table = rc_malloc( sizeof( uint32_t ) * 65536 );
	rc_build_screenmap( table, rmask, gmask, bmask );
	int rcode = rc_dev_set_screenmap( table, (int)bpp );
	rc_free( table );
	if( rcode )
		return rcode;
	rc.map_valid = 1; rc.map_bpp = bpp; rc.map_r = rmask; rc.map_g = gmask; rc.map_b = bmask;
	return 0;
}

static void rc_blit_screen( void )
{
	void *dst = gEngfuncs.SW_LockBuffer();
	if( !dst )
		return;
	RC_CHECK_DEV( rc_set_display_format( rc.sw_bpp, rc.sw_r, rc.sw_g, rc.sw_b ));
	if( rc.nviews == 1 )
	{
		RC_CHECK_DEV( rc_dev_blit( 0, 1, dst, rc.sw_stride, NULL ));
	}
	else
	{
		/* several passes in one frame share the screen in the reference; composite their viewports */
		const int w = gpGlobals->width, h = gpGlobals->height;
		const size_t px = (size_t)w * h, bpp = rc.sw_bpp;
		byte *all = rc_malloc( px * bpp * rc.nviews );
		RC_CHECK_DEV( rc_dev_blit( 0, rc.nviews, all, w, NULL ));
		for( int i = 0; i < rc.nviews; i++ )
		{
			const int *vr = rc.views[i].vrect;
			for( int y = vr[1]; y < vr[1] + vr[3] && y < h; y++ )
				memcpy((byte *)dst + ((size_t)y * rc.sw_stride + vr[0] ) * bpp, all + ( px * i + (size_t)y * w + vr[0] ) * bpp, (size_t)Q_min( vr[2], w - vr[0] ) * bpp );
		}
		rc_free( all );
	}
	gEngfuncs.SW_UnlockBuffer();
}

static void R_EndFrame( void )
{
	if( rc.nviews > 0 && rc.scene_valid )
	{
		int w = gpGlobals->width, h = gpGlobals->height;
		rc_prepare_batch();
		rc_build_glow();
		/* the frames stay in HBM; the display conversion reads them there */
		RC_CHECK_DEV( rc_dev_render( rc.views, rc.nviews, w, h, 0, NULL, NULL ));
		rc_flush_2d( rc.nviews - 1 ); /* the passes share the screen in the reference: the HUD lands on the last one */
		rc_blit_screen();
		rc.screen_view = rc.nviews - 1; rc.screen_valid = 1;
	}
	rc.nviews = 0; rc.ncmds2d = 0;
}

/* ---- batch extension (include/ref_cuda.h) ---- */

static int rc_render_views( const ref_viewpass_t *views, int nviews, unsigned short *color_out, short *z_out, int clear, qboolean async )
{
	if( !rc.initialized || nviews <= 0 )
		return -1;
	rc_dev_h2d_bytes( 1 );
	rc_prepare_batch();
	rc.nviews = 0;
	rc.batch_wvp = rc.nglow > 0 || rc.ntracers > 0;
	for( int i = 0; i < nviews; i++ )
		rc_queue_view( &views[i] );
	rc_build_glow();
	int r = ( async ? rc_dev_render_async : rc_dev_render )( rc.views, rc.nviews, gpGlobals->width, gpGlobals->height, clear, color_out, z_out );
	rc.nviews = 0;
	rc.last_h2d = rc_dev_h2d_bytes( 0 );
	if( r != 0 )
		gEngfuncs.Host_Error( "ref_cuda: rc_dev_render failed: %s\n", rc_dev_last_error());
	return r;
}

int RefCuda_RenderViews( const ref_viewpass_t *views, int nviews, unsigned short *color_out, short *z_out, int clear )
{
	return rc_render_views( views, nviews, color_out, z_out, clear, false );
}

int RefCuda_RenderViewsAsync( const ref_viewpass_t *views, int nviews, unsigned short *color_out, short *z_out, int clear )
{
	return rc_render_views( views, nviews, color_out, z_out, clear, true );
}

int RefCuda_Wait( void )
{
	if( !rc.initialized )
		return -1;
	if( rc_dev_wait() != 0 )
		gEngfuncs.Host_Error( "ref_cuda: rc_dev_wait failed: %s\n", rc_dev_last_error());
	return 0;
}

int RefCuda_BenchResident( int iters, int flags, float *ms_total, float *ms_stage )
{
	return rc_dev_bench_resident( iters, flags, ms_total, ms_stage );
}

int RefCuda_ReadSurfaceCache( int surf, int mip, unsigned short *out, int maxtexels, int *width, int *height, int *valid )
{
	return rc_dev_read_cache_block( surf, mip, out, maxtexels, width, height, valid );
}

int RefCuda_BlitViews( int first, int nviews, void *dst, unsigned int stride, unsigned int bpp, unsigned int rmask, unsigned int gmask, unsigned int bmask, float *ms_kernel )
{
	if( rc_set_display_format( bpp, rmask, gmask, bmask ))
		return -1;
	return rc_dev_blit( first, nviews, dst, stride, ms_kernel );
}

/* see include/ref_cuda.h: the display conversion over the library's static device scratch - no cudaMalloc, no host malloc */
int RefCuda_BlitPixelsNoAlloc( const unsigned short *pixels, unsigned int npixels, void *dst, unsigned int bpp, unsigned int rmask, unsigned int gmask, unsigned int bmask )
{
	static uint32_t table[65536]; /* the call's screen map (256 KB, static: nothing is allocated on the host either) */
	if( !pixels || !dst || ( bpp != 2 && bpp != 4 ))
	{
		gEngfuncs.Con_Printf( S_ERROR "ref_cuda: RefCuda_BlitPixelsNoAlloc: bad arguments (bpp must be 2 or 4)\n" );
		return -1;
	}
	if( npixels > RC_STATIC_BLIT_MAXPX )
	{
		gEngfuncs.Con_Printf( S_ERROR "ref_cuda: RefCuda_BlitPixelsNoAlloc: %u pixels requested, the static scratch holds at most %d (one 1920x1080 frame)\n", npixels, RC_STATIC_BLIT_MAXPX );
		return -2;
	}
	rc_build_screenmap( table, rmask, gmask, bmask );
	if( rc_dev_blit_static( table, (int)bpp, pixels, npixels, dst ) != 0 )
	{
		gEngfuncs.Con_Printf( S_ERROR "ref_cuda: RefCuda_BlitPixelsNoAlloc: %s\n", rc_dev_last_error());
		return -1;
	}
	return 0;
}

int RefCuda_GetCounters( unsigned long long *out, int n ) { return rc_dev_read_counters((uint64_t *)out, n ); }
unsigned long long RefCuda_LaunchCount( void ) { return rc_dev_launch_count(); }
unsigned long long RefCuda_BatchH2DBytes( void ) { return rc.last_h2d; }
const char *RefCuda_DeviceName( void ) { return rc_dev_name(); }

/* ---------------------------------------------------------------------------------------------- */
/* init / shutdown                                                                                  */

static qboolean R_Init( void )
{
	gEngfuncs.Cvar_RegisterVariable( &sw_clearcolor );
	gEngfuncs.Cvar_RegisterVariable( &sw_mipcap );
	gEngfuncs.Cvar_RegisterVariable( &sw_texfilt );
	gEngfuncs.Cvar_RegisterVariable( &sw_mipscale );
	gEngfuncs.Cvar_RegisterVariable( &sw_noalphabrushes );
	gEngfuncs.Cvar_RegisterVariable( &cu_device );
	gEngfuncs.Cvar_RegisterVariable( &r_novis );
	gEngfuncs.Cvar_RegisterVariable( &r_traceglow );
	gEngfuncs.Cvar_RegisterVariable( &r_dlight_virtual_radius );

	rc.pool = gEngfuncs._Mem_AllocPool( "ref_cuda zone", 0, __FILE__, __LINE__ );

	if( !gEngfuncs.R_Init_Video( REF_SOFTWARE ))
	{
		gEngfuncs.R_Free_Video();
		gEngfuncs.Con_Printf( S_ERROR "ref_cuda: can't initialize video subsystem\n" );
		return false;
	}

	/* no CPU path exists behind this plugin: without a usable CUDA device initialisation fails */
	if( rc_dev_init((int)cu_device.value ) != 0 )
	{
		gEngfuncs.Con_Printf( S_ERROR "ref_cuda: CUDA initialisation failed: %s\n", rc_dev_last_error());
		gEngfuncs.R_Free_Video();
		return false;
	}

	rc_resolve_lightgamma();
	memcpy( rc.lightgamma_sent, rc.lightgamma, sizeof( rc.lightgamma_sent ));

	rc_build_colormap();
	for( int i = 0; i < 1280; i++ ) /* R_InitTurb, ref/soft/r_main.c:1312: AMP + sin( i * 3.14159 * 2 / CYCLE ) * AMP */
		rc.sintable[i] = ( 8 * 0x10000 ) + sin( i * 3.14159 * 2 / 128 ) * ( 8 * 0x10000 );
	RC_CHECK_DEV( rc_dev_upload_luts( rc.colormap, rc.lightgamma, rc.sintable ));
	{
		uint16_t *alphamap = rc_malloc( sizeof( uint16_t ) * 3 * 1024 * 256 );
		uint8_t *addmap = rc_malloc( 65536 ), *modmap = rc_malloc( 65536 );
		rc_build_blendmaps( alphamap, addmap, modmap );
		RC_CHECK_DEV( rc_dev_upload_blendluts( alphamap, addmap, modmap ));
		rc_free( alphamap ); rc_free( addmap ); rc_free( modmap );
	}
	rc.tri.color = 0xFFFF; rc.tri.alpha = 7; rc.tri.blend = 1.0f;

	if( gpGlobals->width < 128 ) gpGlobals->width = 128;
	if( gpGlobals->height < 128 ) gpGlobals->height = 128;
	if( !gEngfuncs.SW_CreateBuffer( gpGlobals->width, gpGlobals->height, &rc.sw_stride, &rc.sw_bpp, &rc.sw_r, &rc.sw_g, &rc.sw_b ))
		return false;

	/* texture slot 0 is never handed out (r_image.c:859) */
	memset( rc.images, 0, sizeof( rc.images ));
	rc_strncpy( rc.images[0].name, "*unused*", sizeof( rc.images[0].name ));
	rc.numImages = 1;

	/* GL_InitRandomTable, ref/common/ref_context.c:46-56 */
	for( int tu = 0; tu < MOD_FRAMES; tu++ )
		for( int tv = 0; tv < MOD_FRAMES; tv++ )
			rc_tile_random[tu][tv] = gEngfuncs.COM_RandomLong( 0, 0x7FFF );
	gEngfuncs.COM_SetRandomSeed( 0 );

	for( int i = 0; i < MAX_LIGHTSTYLES; i++ ) rc.lightstylevalue[i] = 256;
	rc.initialized = true;
	gEngfuncs.Con_Printf( "ref_cuda: rendering on %s\n", rc_dev_name());
	return true;
}

static void R_Shutdown( void )
{
	if( !rc.initialized )
		return;
	rc_dev_shutdown();
	gEngfuncs.R_Free_Video();
	gEngfuncs._Mem_FreePool( &rc.pool, __FILE__, __LINE__ );
	memset( &rc, 0, sizeof( rc ));
}

static const char *R_GetConfigName( void ) { return "ref_cuda"; }

/* ---------------------------------------------------------------------------------------------- */
/* the remaining ref_interface_t slots.  Every slot is non-NULL (the engine calls most of them
 * unconditionally); the ones ref_soft itself leaves empty (ref/soft/r_context.c:23-26,266-375) are
 * empty here too.  Entity, studio, sprite, decal and 2D slots accept their calls and are wired to
 * the GPU path as those stages land. */

static qboolean R_SetDisplayTransform( ref_screen_rotation_t rotate, int x, int y, float sx, float sy )
{
	(void)x; (void)y; (void)sx; (void)sy;
	return rotate == REF_ROTATE_NONE;
}
static void GL_SetupAttributes( int safegl ) { (void)safegl; }
static void GL_InitExtensions( void ) {}
static void GL_ClearExtensions( void ) {}
static void R_GammaChanged( qboolean do_reset_gamma ) { if( !do_reset_gamma && rc.initialized ) { rc_sync_lightgamma(); rc_dev_flush_cache(); } }
static void R_PushScene( void ) {}
static void R_PopScene( void ) {}
static void GL_BackendStartFrame( void ) {}
static void GL_BackendEndFrame( void ) {}
static void R_ClearScreen( void ) {}
static void R_AllowFog( qboolean allow ) { (void)allow; }
static void R_ShowTextures( void ) {}
static const byte *R_GetTextureOriginalBuffer( unsigned int idx ) { (void)idx; return NULL; }
static void GL_ProcessTexture( int texnum, float gamma, int topColor, int bottomColor ) { (void)texnum; (void)gamma; (void)topColor; (void)bottomColor; }
static void R_SetupSky( int *skyboxTextures ) { (void)skyboxTextures; }
/* R_Set2DMode (ref/soft/r_draw.c:228) */
static void R_Set2DMode( qboolean enable )
{
	rc.tri.color = 0xFFFF; /* COLOR_WHITE */
	rc.tri.is2d = enable;
	rc.tri.alpha = 7;
}

static rc_cmd2d_t *rc_cmd2d_new( void )
{
	if( rc.ncmds2d == rc.maxcmds2d )
	{
		int n = rc.maxcmds2d ? rc.maxcmds2d * 2 : 64;
		rc.cmds2d = gEngfuncs._Mem_Realloc( rc.pool, rc.cmds2d, sizeof( rc_cmd2d_t ) * n, true, __FILE__, __LINE__ );
		rc.maxcmds2d = n;
	}
	rc_cmd2d_t *c = &rc.cmds2d[rc.ncmds2d++];
	memset( c, 0, sizeof( *c ));
	c->color = rc.tri.color; c->alpha = rc.tri.alpha; c->rendermode = rc.tri.rendermode;
	return c;
}

/* R_DrawStretchPic + the clipping half of R_DrawStretchPicImplementation (ref/soft/r_draw.c:38-85,145): the HUD pass is
 * recorded and applied to the frame on the device once the views are drawn (R_EndFrame), in call order */
static void R_DrawStretchPic( float fx, float fy, float fw, float fh, float fs1, float ft1, float fs2, float ft2, int texnum )
{
	if( texnum < 0 || texnum >= RC_MAX_TEXTURES ) texnum = 0;
	const rc_image_t *pic = &rc.images[texnum];
	if( fs2 > 1.0f || ft2 > 1.0f || fs1 < 0.0f || ft1 < 0.0f || fw < 1.0f || fh < 1.0f )
		return;
	int x = fx, y = fy, w = fw, h = fh;
	int s1 = pic->width * fs1, t1 = pic->height * ft1, s2 = pic->width * fs2, t2 = pic->height * ft2;
	const int vw = gpGlobals->width, vh = gpGlobals->height;
	if( x < 0 ) { s1 += ( -x ) * ( s2 - s1 ) / w; x = 0; }
	if( x + w > vw ) { s2 -= ( x + w - vw ) * ( s2 - s1 ) / w; w = vw - x; }
	if( y + h > vh ) { t2 -= ( y + h - vh ) * ( t2 - t1 ) / h; h = vh - y; }
	if( !RC_IMAGE_LIVE( pic ) || s1 >= s2 || t1 >= t2 )
		return;
	unsigned int height = h;
	int skip = 0;
	if( y < -h )
		return;
	if( y < 0 ) { skip = -y; height += y; y = 0; }
	if( w <= 0 || h <= 0 || (int)height <= 0 )
		return; /* the reference's loops do not run */
	rc_cmd2d_t *c = rc_cmd2d_new();
	c->kind = RC_2D_PIC; c->x = x; c->y = y; c->w = w; c->h = h; c->s1 = s1; c->t1 = t1; c->s2 = s2; c->t2 = t2;
	c->skip = skip; c->height = (int)height;
	c->tex_ofs = texnum; /* resolved to an arena offset when the pass is sent */
	c->texwidth = pic->width; c->transparent = pic->has_alpha;
}

/* CL_FillRGBA + Draw_Fill (ref/soft/r_context.c:44, r_draw.c:159) */
static void FillRGBA( int rendermode, float fx, float fy, float fw, float fh, byte r, byte g, byte b, byte a )
{
	rc.tri.rendermode = rendermode;
	rc_tri_color4f( r * ( 1.0f / 255.0f ), g * ( 1.0f / 255.0f ), b * ( 1.0f / 255.0f ), a * ( 1.0f / 255.0f ));
	int x = fx, y = fy, w = fw, h = fh;
	const int vw = gpGlobals->width, vh = gpGlobals->height;
	if( x < 0 ) x = 0;
	if( x + w > vw ) w = vw - x;
	if( w <= 0 ) return;
	if( y + h > vh ) h = vh - y;
	if( h <= 0 ) return;
	unsigned int height = h;
	if( y < 0 )
	{
		if( h <= -y ) return;
		height += y; y = 0;
	}
	rc_cmd2d_t *c = rc_cmd2d_new();
	c->kind = RC_2D_FILL; c->x = x; c->y = y; c->w = w; c->h = h; c->height = (int)height; c->tex_ofs = -1;
}

/* the recorded HUD pass onto frame `view` of the batch in HBM */
static void rc_flush_2d( int view )
{
	int n = 0;
	for( int i = 0; i < rc.ncmds2d; i++ )
	{
		rc_cmd2d_t c = rc.cmds2d[i];
		if( c.kind == RC_2D_PIC )
		{
			const rc_texture_t *t = rc_texture_entry( c.tex_ofs );
			if( !t ) continue;
			c.tex_ofs = c.transparent ? t->alpha_ofs : t->ofs[0];
			if( c.tex_ofs < 0 ) continue;
		}
		rc.cmds2d[n++] = c;
	}
	if( n ) RC_CHECK_DEV( rc_dev_draw2d( rc.cmds2d, n, view ));
	rc.ncmds2d = 0;
}

/* R_WorldToScreen (ref/soft/r_main.c:210) against the matrix of the view set up last, as RI keeps it between frames */
static int WorldToScreen( const vec3_t world, vec3_t screen )
{
	if( !world || !screen )
		return true;
	rc_view_t v;
	rc_view_wvp( &rc.last_rvp, &v );
	return rc_world_to_screen_m( v.wvp, world, screen );
}
/* VID_ScreenShot (ref/soft/r_glblit.c:748): the frame last shown, in display colours, bottom row first, as PF_BGRA_32
 * through the engine's image writer.  The frame is still in HBM: k_blit converts it once more, into the shot. */
static qboolean VID_ScreenShot( const char *filename, int shot_type )
{
	const int vw = gpGlobals->width, vh = gpGlobals->height;
	if( !rc.screen_valid )
		return false;
	uint flags = IMAGE_FLIP_Y;
	int width = 0, height = 0;
	rgbdata_t *r_shot = gEngfuncs._Mem_Alloc( rc.pool, sizeof( rgbdata_t ), true, __FILE__, __LINE__ );
	r_shot->width = ( vw + 3 ) & ~3;
	r_shot->height = ( vh + 3 ) & ~3;
	r_shot->flags = IMAGE_HAS_COLOR;
	r_shot->type = PF_BGRA_32;
	r_shot->size = r_shot->width * r_shot->height * gEngfuncs.Image_GetPFDesc( r_shot->type )->bpp;
	r_shot->palette = NULL;
	r_shot->buffer = gEngfuncs._Mem_Alloc( rc.pool, r_shot->size, false, __FILE__, __LINE__ );

	byte *disp = rc_malloc((size_t)vw * vh * rc.sw_bpp );
	RC_CHECK_DEV( rc_set_display_format( rc.sw_bpp, rc.sw_r, rc.sw_g, rc.sw_b ));
	RC_CHECK_DEV( rc_dev_blit( rc.screen_view, 1, disp, vw, NULL ));
	uint32_t *pbuf = (uint32_t *)r_shot->buffer;
	for( int v = 0; v < vh; v++ )
	{
		const size_t start = (size_t)vw * ( vh - v - 1 ), dstart = (size_t)rc.sw_stride * v;
		for( int u = 0; u < vw; u++ )
		{
			uint32_t s32;
			if( rc.sw_bpp == 2 )
			{
				const uint color = ((const uint16_t *)disp )[start + u];
				const uint c0 = (((( color >> 11 ) & 0x1F ) * 527 ) + 23 ) >> 6, c1 = (((( color >> 5 ) & 0x3F ) * 259 ) + 33 ) >> 6, c2 = ((( color & 0x1F ) * 527 ) + 23 ) >> 6;
				s32 = ( c0 & 0xFF ) << 16 | ( c1 & 0xFF ) << 8 | ( c2 & 0xFF );
			}
			else
				s32 = ((const uint32_t *)disp )[start + u];
			pbuf[dstart + u] = s32 | 0xFF000000u;
		}
	}
	rc_free( disp );

	switch( shot_type )
	{
	case VID_SCREENSHOT: break;
	case VID_SNAPSHOT: gEngfuncs.fsapi->AllowDirectPaths( true ); break;
	case VID_LEVELSHOT:
	case VID_MINISHOT:
		flags |= IMAGE_RESAMPLE;
		height = shot_type == VID_MINISHOT ? 200 : 480;
		width = Q_rint( height * ((double)r_shot->width / r_shot->height ));
		break;
	case VID_MAPSHOT:
		flags |= IMAGE_RESAMPLE | IMAGE_QUANTIZE;
		height = 768; width = 1024;
		break;
	}
	gEngfuncs.Image_Process( &r_shot, width, height, flags, 0.0f );
	qboolean result = gEngfuncs.FS_SaveImage( filename, r_shot );
	gEngfuncs.fsapi->AllowDirectPaths( false );
	gEngfuncs.FS_FreeImage( r_shot );
	return result;
}
static qboolean VID_CubemapShot( const char *base, uint size, const float *vieworg, qboolean skyshot ) { (void)base; (void)size; (void)vieworg; (void)skyshot; return false; }
/* R_LightPoint (ref/common/ref_light.c:471): the light the world throws on the ground below p */
static colorVec R_LightPoint( const float *p )
{
	const vec3_t p0 = { p[0], p[1], p[2] }, p1 = { p[0], p[1], p[2] - 2048.0f };
	uint light[3];
	rc_light_vec( p0, p1, false, light );
	colorVec c = { light[0], light[1], light[2], 255 };
	return c;
}
/* ---------------------------------------------------------------------------------------------- */
/*
 * N2 decals: placement (ref/soft/r_decals.c) on the host, drawing into the surface cache (R_DrawSurfaceDecals,
 * r_surf.c:914) on the device.  A decal is a decal_t of the engine's ABI hung on msurface_t::pdecals, as in every
 * renderer of the tree (the engine's save / restore walks those lists).
 *
 * ref_soft keeps a new decal only when R_DecalVertsClip reports that the face's polygon, clipped to the decal's unit
 * square in decal texture space, has vertices left (r_decals.c:356, :578).  It reads the polygon from msurface_t::polys,
 * which only the GL renderers build: in ref_soft the pointer is NULL, the function returns before it writes the count
 * and R_DecalCreate tests an uninitialised local.  ref/cuda takes the designed path - it clips the polygon it reads from
 * the BSP's own edges - and the test harness gives the oracle's faces their polygons so that the reference takes it too.
 */
#define RC_DECAL_DISTANCE   4
#define RC_MAX_OVERLAP      6     /* MAX_OVERLAP_DECALS */
#define RC_DECAL_CLIPVERTS  32    /* MAX_DECALCLIPVERT */

static decal_t rc_decal_pool[MAX_RENDER_DECALS];
static int     rc_decal_next;
static qboolean rc_decals_dirty; /* the device's decal lists are out of date */
static uint32_t *rc_decal_gen;   /* per world face: bumped when a decal is hung on it (forces the rebuild of its cache blocks) */
static int      rc_decal_gen_n;

typedef struct
{
	vec3_t  position;
	model_t *model;
	int     texture, size, flags, entity;
	float   scale;
	int     width, height;   /* world size of the picture */
	vec3_t  basis[3];
} rc_decalshot_t;

static void rc_decals_texels_moved( void ) { rc_decals_dirty = true; }

static void rc_decal_gen_ensure( void )
{
	if( rc.world && rc_decal_gen_n != rc.world->numsurfaces )
	{
		rc_free( rc_decal_gen );
		rc_decal_gen_n = rc.world->numsurfaces;
		rc_decal_gen = rc_malloc( sizeof( uint32_t ) * Q_max( 1, rc_decal_gen_n ));
		memset( rc_decal_gen, 0, sizeof( uint32_t ) * Q_max( 1, rc_decal_gen_n ));
	}
}

static void rc_decal_gen_bump( msurface_t *surf )
{
	if( !rc.world )
		return;
	rc_decal_gen_ensure();
	const int i = (int)( surf - rc.world->surfaces );
	if( i >= 0 && i < rc_decal_gen_n )
		rc_decal_gen[i]++;
	rc_decals_dirty = true;
}

static void rc_decal_unlink( decal_t *d )
{
	msurface_t *surf = d->psurface;
	if( surf )
	{
		decal_t **link = &surf->pdecals;
		while( *link && *link != d )
			link = &( *link )->pnext;
		if( *link == d )
			*link = d->pnext;
		rc_decals_dirty = true; /* the lists changed; the face's cached blocks stay as they are, as in the reference */
	}
	d->psurface = NULL;
	d->polys = NULL;
}

/* R_DecalAlloc: the entry to overwrite - the one handed in, or the next of the ring that is not permanent */
static decal_t *rc_decal_alloc( decal_t *d )
{
	int limit = MAX_RENDER_DECALS;
	if( r_decals->value < limit )
		limit = r_decals->value;
	if( !limit )
		return NULL;
	if( !d )
	{
		int count = 0;
		do
		{
			if( rc_decal_next >= limit )
				rc_decal_next = 0;
			d = &rc_decal_pool[rc_decal_next++];
			count++;
		} while( FBitSet( d->flags, FDECAL_PERMANENT ) && count < limit );
	}
	rc_decal_unlink( d );
	return d;
}

/* R_DecalComputeBasis: the face's texture axes and normal, unit length */
static void rc_decal_basis( const msurface_t *surf, vec3_t basis[3] )
{
	vec3_t n;
	if( surf->flags & SURF_PLANEBACK ) VectorNegate( surf->plane->normal, n );
	else VectorCopy( surf->plane->normal, n );
	VectorNormalize2( n, basis[2] );
	VectorNormalize2( surf->texinfo->vecs[0], basis[0] );
	VectorNormalize2( surf->texinfo->vecs[1], basis[1] );
}

/* R_SetupDecalTextureSpaceBasis: world -> decal texture space ([0,1] over the picture) */
static void rc_decal_texbasis( const decal_t *d, const msurface_t *surf, int texture, vec3_t basis[3], float worldscale[2] )
{
	rc_decal_basis( surf, basis );
	const rc_image_t *img = &rc.images[( texture >= 0 && texture < RC_MAX_TEXTURES ) ? texture : 0];
	int width = 1, height = 1; /* R_GetDecalDimensions */
	width = img->srcWidth; height = img->srcHeight;
	worldscale[0] = (float)d->scale / width;
	worldscale[1] = (float)d->scale / height;
	VectorScale( basis[0], worldscale[0], basis[0] );
	VectorScale( basis[1], worldscale[1], basis[1] );
}

/* one pass of the Sutherland-Hodgman clip (SHClip, r_decals.c:292) over decal-space (u, v) pairs */
static int rc_decal_clip_edge( const float ( *in )[2], int n, float ( *out )[2], int edge )
{
	int m = 0;
	const int axis = edge >> 1;            /* 0: u (left, right), 1: v (top, bottom) */
	const float lim = ( edge & 1 ) ? 1.0f : 0.0f;
#define RC_INSIDE( p ) (( edge & 1 ) ? ( p )[axis] < 1.0f : ( p )[axis] > 0.0f )
	const float *s = in[n - 1];
	for( int j = 0; j < n; j++ )
	{
		const float *p = in[j];
		const qboolean pin = RC_INSIDE( p ), sin_ = RC_INSIDE( s );
		if( pin != sin_ )
		{
			/* R_ClipIntersect( one, two ): entering: one = s, two = p; leaving: one = p, two = s */
			const float *one = pin ? s : p, *two = pin ? p : s;
			const float t = (( one[axis] - lim ) / ( one[axis] - two[axis] ));
			if( m < RC_DECAL_CLIPVERTS )
			{
				out[m][axis] = lim;
				out[m][axis ^ 1] = one[axis ^ 1] + ( two[axis ^ 1] - one[axis ^ 1] ) * t;
				m++;
			}
		}
		if( pin && m < RC_DECAL_CLIPVERTS )
		{
			out[m][0] = p[0]; out[m][1] = p[1];
			m++;
		}
		s = p;
	}
#undef RC_INSIDE
	return m;
}

/* R_DecalVertsClip: sets the decal's dx, dy (R_SetupDecalClip) and counts what is left of the face's polygon inside
 * the decal's square */
static int rc_decal_clip_face( decal_t *d, const msurface_t *surf, const model_t *mod, int texture )
{
	vec3_t basis[3];
	float worldscale[2];
	rc_decal_texbasis( d, surf, texture, basis, worldscale );
	d->dx = DotProduct( d->position, basis[0] );
	d->dy = DotProduct( d->position, basis[1] );
	float a[RC_DECAL_CLIPVERTS][2], b[RC_DECAL_CLIPVERTS][2];
	int n = 0;
	for( int i = 0; i < surf->numedges && n < RC_DECAL_CLIPVERTS; i++ )
	{
		const int e = mod->surfedges[surf->firstedge + i];
		const float *v;
		if( FBitSet( mod->flags, MODEL_QBSP2 )) v = e >= 0 ? mod->vertexes[mod->edges32[e].v[0]].position : mod->vertexes[mod->edges32[-e].v[1]].position;
		else v = e >= 0 ? mod->vertexes[mod->edges16[e].v[0]].position : mod->vertexes[mod->edges16[-e].v[1]].position;
		a[n][0] = DotProduct( v, basis[0] ) - d->dx + 0.5f;
		a[n][1] = DotProduct( v, basis[1] ) - d->dy + 0.5f;
		n++;
	}
	if( !n )
		return 0;
	n = rc_decal_clip_edge( a, n, b, 0 );
	if( n ) n = rc_decal_clip_edge( b, n, a, 1 );
	if( n ) n = rc_decal_clip_edge( a, n, b, 2 );
	if( n ) n = rc_decal_clip_edge( b, n, a, 3 );
	return n;
}

/* R_DecalIntersect: how many decals of the face the new one would mostly cover, and the smallest of them */
static decal_t *rc_decal_intersect( const rc_decalshot_t *shot, msurface_t *surf, int *pcount )
{
	decal_t *last = NULL;
	float lastarea = 2;
	vec3_t ext[2];
	const rc_image_t *img = &rc.images[shot->texture];
	*pcount = 0;
	VectorScale( shot->basis[0], (( img->srcWidth / shot->scale ) * 0.5f ), ext[0] );
	VectorScale( shot->basis[1], (( img->srcHeight / shot->scale ) * 0.5f ), ext[1] );
	for( decal_t *d = surf->pdecals; d; d = d->pnext )
	{
		if( !FBitSet( d->flags, FDECAL_PERMANENT ))
		{
			vec3_t tb[3], tp[2];
			float ws[2];
			rc_decal_texbasis( d, surf, d->texture, tb, ws );
			VectorSubtract( shot->position, ext[0], tp[0] );
			VectorSubtract( shot->position, ext[1], tp[1] );
			const float minx = DotProduct( tp[0], tb[0] ) - d->dx + 0.5f, miny = DotProduct( tp[1], tb[1] ) - d->dy + 0.5f;
			VectorAdd( shot->position, ext[0], tp[0] );
			VectorAdd( shot->position, ext[1], tp[1] );
			const float maxx = DotProduct( tp[0], tb[0] ) - d->dx + 0.5f, maxy = DotProduct( tp[1], tb[1] ) - d->dy + 0.5f;
			const float umin[2] = { Q_max( minx, 0 ), Q_max( miny, 0 ) }, umax[2] = { Q_min( maxx, 1 ), Q_min( maxy, 1 ) };
			if( umin[0] < 1 && umin[1] < 1 && umax[0] > 0 && umax[1] > 0 )
			{
				/* the reference's expression, umin[1] in the first factor included */
				const float area = ( umax[0] - umin[1] ) * ( umax[1] - umin[1] );
				if( area > 0.6f )
				{
					*pcount += 1;
					if( !last || area <= lastarea ) { last = d; lastarea = area; }
				}
			}
		}
		if( d == d->pnext ) break;
	}
	return last;
}

/* R_DecalCreate + R_AddDecalToSurface */
static void rc_decal_create( const rc_decalshot_t *shot, msurface_t *surf, float x, float y )
{
	int count;
	decal_t *old = rc_decal_intersect( shot, surf, &count );
	if( count < RC_MAX_OVERLAP )
		old = NULL;
	decal_t *d = rc_decal_alloc( old );
	if( !d )
		return;
	d->flags = shot->flags;
	VectorCopy( shot->position, d->position );
	d->dx = x; d->dy = y;
	d->scale = shot->scale;
	d->entityIndex = shot->entity;
	d->texture = shot->texture;
	if( !rc_decal_clip_face( d, surf, shot->model, shot->texture ))
	{
		rc_decal_unlink( d );
		return;
	}
	d->pnext = NULL;
	decal_t *tail = surf->pdecals;
	if( tail )
	{
		while( tail->pnext && tail != tail->pnext )
			tail = tail->pnext;
		tail->pnext = d;
	}
	else surf->pdecals = d;
	d->psurface = surf;
	rc_decal_gen_bump( surf ); /* surf->dlightframe = tr.framecount + 1: every cached block of the face is rebuilt */
}

/* R_DecalSurface: does the picture's rectangle touch the face */
static void rc_decal_surface( msurface_t *surf, rc_decalshot_t *shot )
{
	const mtexinfo_t *tex = surf->texinfo;
	const connstate_t state = (connstate_t)gEngfuncs.EngineGetParm( PARM_CONNSTATE, 0 );
	if( state == ca_connected || state == ca_validate )
	{
		/* restoring a level: a decal that came over from the neighbouring one is not placed twice */
		for( decal_t *d = surf->pdecals; d && d != d->pnext; d = d->pnext )
			if( VectorCompare( d->position, shot->position ) && d->texture == shot->texture )
				return;
	}
	float s = DotProduct( shot->position, tex->vecs[0] ) + tex->vecs[0][3] - surf->texturemins[0];
	float t = DotProduct( shot->position, tex->vecs[1] ) + tex->vecs[1][3] - surf->texturemins[1];
	rc_decal_basis( surf, shot->basis );
	const float w = fabs( shot->width * DotProduct( tex->vecs[0], shot->basis[0] )) + fabs( shot->height * DotProduct( tex->vecs[0], shot->basis[1] ));
	const float h = fabs( shot->width * DotProduct( tex->vecs[1], shot->basis[0] )) + fabs( shot->height * DotProduct( tex->vecs[1], shot->basis[1] ));
	s -= ( w * 0.5f );
	t -= ( h * 0.5f );
	if( s <= -w || t <= -h || s > ( surf->extents[0] + w ) || t > ( surf->extents[1] + h ))
		return;
	rc_decal_create( shot, surf, s, t );
}

/* R_DecalNode: the nodes whose plane passes within the picture's size of the point */
static void rc_decal_node( model_t *model, mnode_t *node, rc_decalshot_t *shot )
{
	while( node->contents >= 0 )
	{
		const mplane_t *plane = node->plane;
		const float dist = DotProduct( shot->position, plane->normal ) - plane->dist;
		if( dist > shot->size ) { node = node_child( node, 0, model ); continue; }
		if( dist < -shot->size ) { node = node_child( node, 1, model ); continue; }
		if( dist < RC_DECAL_DISTANCE && dist > -RC_DECAL_DISTANCE )
		{
			msurface_t *surf = model->surfaces + node_firstsurface( node, model );
			for( int i = 0, n = node_numsurfaces( node, model ); i < n; i++, surf++ )
				if( !( surf->flags & ( SURF_DRAWTURB | SURF_DRAWSKY | SURF_CONVEYOR )))
					rc_decal_surface( surf, shot );
		}
		rc_decal_node( model, node_child( node, 0, model ), shot );
		node = node_child( node, 1, model );
	}
}

/* CL_ModelHandle (r_local.h:505) */
static model_t *rc_model_handle( int index )
{
	return ( index >= 0 && index < gp_cl->nummodels ) ? gp_cl->models[index] : NULL;
}

/* R_DecalShoot (ref/soft/r_decals.c:719) */
static void R_DecalShoot( int textureIndex, int entityIndex, int modelIndex, vec3_t pos, int flags, float scale )
{
	rc_decalshot_t shot;
	cl_entity_t *ent = NULL;
	model_t *model = NULL;
	if( textureIndex <= 0 || textureIndex >= RC_MAX_TEXTURES )
	{
		gEngfuncs.Con_Printf( S_ERROR "Decal has invalid texture!\n" );
		return;
	}
	if( entityIndex > 0 )
	{
		ent = ( rc.entities && (uint)entityIndex < rc.max_entities ) ? &rc.entities[entityIndex] : NULL;
		if( modelIndex > 0 ) model = rc_model_handle( modelIndex );
		else if( ent != NULL ) model = rc_model_handle( ent->curstate.modelindex );
		else return;
	}
	else if( modelIndex > 0 ) model = rc_model_handle( modelIndex );
	else model = gp_cl->models[1];
	if( !model )
		return;
	if( model->type != mod_brush )
	{
		gEngfuncs.Con_Reportf( S_ERROR "Decals must hit mod_brush!\n" );
		return;
	}
	memset( &shot, 0, sizeof( shot ));
	shot.model = model;
	const hull_t *hull = &model->hulls[0];
	if( ent && !FBitSet( flags, FDECAL_LOCAL_SPACE ))
	{
		/* into the brush entity's own space */
		vec3_t local;
		if( !VectorIsNull( ent->angles ))
		{
			matrix4x4 m;
			Matrix4x4_CreateFromEntity( m, ent->angles, ent->origin, 1.0f );
			Matrix4x4_VectorITransform( m, pos, local );
		}
		else VectorSubtract( pos, ent->origin, local );
		VectorCopy( local, shot.position );
		SetBits( flags, FDECAL_LOCAL_SPACE );
	}
	else VectorCopy( pos, shot.position );
	if( !FBitSet( model->flags, MODEL_HAS_ORIGIN ))
		SetBits( flags, FDECAL_USE_LANDMARK );
	shot.texture = textureIndex;
	shot.entity = entityIndex;
	shot.flags = flags;
	const int width = rc.images[textureIndex].srcWidth, height = rc.images[textureIndex].srcHeight;
	shot.size = width >> 1;
	if(( height >> 1 ) > shot.size )
		shot.size = height >> 1;
	shot.scale = bound( 0.01f, scale, 16.0f ); /* MIN_DECAL_SCALE, MAX_DECAL_SCALE */
	shot.width = width / shot.scale;
	shot.height = height / shot.scale;
	rc_decal_node( model, &model->nodes[hull->firstclipnode], &shot );
}

/* R_DecalRemoveAll (r_decals.c:978): the decals of one texture (0: of any) among the pool entries handed out so far */
static void R_DecalRemoveAll( int texture )
{
	if( texture < 0 || texture >= RC_MAX_TEXTURES )
		return;
	for( int i = 0; i < rc_decal_next; i++ )
	{
		decal_t *d = &rc_decal_pool[i];
		if( FBitSet( d->flags, FDECAL_PERMANENT ))
			continue;
		if( !texture || d->texture == texture )
			rc_decal_unlink( d );
	}
}

/* the file's base name: no directories, no extension (COM_FileBase, public/crtlib.c:491) */
static void rc_file_base( const char *in, char *out, size_t size )
{
	out[0] = 0;
	if( !in || !in[0] || size <= 1 )
		return;
	const char *start = in, *dot = NULL, *p;
	for( p = in; *p; p++ )
	{
		if( *p == '/' || *p == '\\' ) start = p + 1;
		if( *p == '.' ) dot = p;
	}
	if( !dot || dot < start )
		dot = p;
	const size_t len = Q_min( size - 1, (size_t)( dot - start ));
	memcpy( out, start, len );
	out[len] = 0;
}

static int rc_decal_depth_compare( const void *a, const void *b )
{
	const decallist_t *x = a, *y = b;
	return x->depth > y->depth ? 1 : ( x->depth < y->depth ? -1 : 0 );
}

/* R_CreateDecalList (r_decals.c:916): the decals of the level for a save game - where, which picture, how deep in its
 * face's list - without the near-duplicates (same picture and entity within 2 units), lowest depth first */
static int R_CreateDecalList( struct decallist_s *pList )
{
	int total = 0;
	if( rc.world )
	{
		for( int i = 0; i < MAX_RENDER_DECALS; i++ )
		{
			const decal_t *d = &rc_decal_pool[i];
			if( d->psurface == NULL || FBitSet( d->flags, FDECAL_DONTSAVE ))
				continue;
			int depth = 0;
			for( const decal_t *q = d->psurface->pdecals; q && q != d; q = q->pnext )
				depth++;
			decallist_t *e = &pList[total];
			e->depth = depth;
			e->flags = d->flags;
			e->scale = d->scale;
			VectorCopy( d->position, e->position );
			e->entityIndex = d->entityIndex;
			if( d->psurface->flags & SURF_PLANEBACK ) VectorNegate( d->psurface->plane->normal, e->impactPlaneNormal );
			else VectorCopy( d->psurface->plane->normal, e->impactPlaneNormal );
			rc_file_base( rc.images[( d->texture >= 0 && d->texture < RC_MAX_TEXTURES ) ? d->texture : 0].name, e->name, sizeof( e->name ));
			/* DecalListAdd: kept unless an entry before it is the same decal again */
			qboolean merged = false;
			for( int k = 0; k < total && !merged; k++ )
				if( !strcmp( e->name, pList[k].name ) && e->entityIndex == pList[k].entityIndex )
				{
					vec3_t tmp;
					VectorSubtract( e->position, pList[k].position, tmp );
					if( VectorLength( tmp ) < 2 ) /* DECAL_OVERLAP_DISTANCE */
						merged = true;
				}
			if( !merged )
				total++;
		}
		if( gEngfuncs.drawFuncs->R_CreateStudioDecalList )
			total += gEngfuncs.drawFuncs->R_CreateStudioDecalList( pList, total );
	}
	qsort( pList, total, sizeof( decallist_t ), rc_decal_depth_compare );
	return total;
}

/* R_ClearAllDecals (r_decals.c:1024): every pool entry leaves its face */
static void R_ClearAllDecals( void )
{
	for( int i = 0; i < MAX_RENDER_DECALS; i++ )
		rc_decal_unlink( &rc_decal_pool[i] );
}

/* R_ClearDecals (r_decals.c:57), at R_NewMap: the pool starts over */
static void rc_clear_decals( void )
{
	memset( rc_decal_pool, 0, sizeof( rc_decal_pool ));
	rc_decal_next = 0;
	rc_free( rc_decal_gen ); rc_decal_gen = NULL; rc_decal_gen_n = 0;
	rc_decals_dirty = true;
}

/*
 * The device's view of the lists: per world face the slice of its decals, per decal and mip level the rectangle of the
 * block it covers and the part of the picture it shows - the integer set-up of R_DrawSurfaceDecals (r_surf.c:914-990),
 * which depends on the face, the decal and the mip only.
 */
static void rc_send_decals( void )
{
	if( !rc_decals_dirty || !rc.world )
		return;
	rc_decals_dirty = false;
	const model_t *w = rc.world;
	const int ns = w->numsurfaces;
	int32_t *first = rc_malloc( sizeof( int32_t ) * 2 * Q_max( 1, ns ));
	int total = 0;
	for( int i = 0; i < ns; i++ )
		for( const decal_t *d = w->surfaces[i].pdecals; d; d = ( d == d->pnext ) ? NULL : d->pnext )
			total++;
	rc_ddecal_t *recs = rc_malloc( sizeof( rc_ddecal_t ) * Q_max( 1, total ));
	memset( recs, 0, sizeof( rc_ddecal_t ) * Q_max( 1, total ));
	int n = 0;
	for( int i = 0; i < ns; i++ )
	{
		const msurface_t *fa = &w->surfaces[i];
		first[2 * i] = n; first[2 * i + 1] = 0;
		for( const decal_t *p = fa->pdecals; p; p = ( p == p->pnext ) ? NULL : p->pnext )
		{
			rc_ddecal_t *o = &recs[n];
			const int texnum = ( p->texture >= 0 && p->texture < RC_MAX_TEXTURES ) ? p->texture : 0;
			const rc_image_t *img = &rc.images[texnum];
			const rc_texture_t *te = rc_texture_entry( texnum );
			if( !RC_IMAGE_LIVE( img ) || !te )
				continue; /* !tex->pixels[0] */
			vec3_t basis[3];
			const float *tu = fa->texinfo->vecs[0], *tv = fa->texinfo->vecs[1];
			rc_decal_basis( fa, basis );
			const int w0 = fabs( img->width * DotProduct( tu, basis[0] )) + fabs( img->height * DotProduct( tu, basis[1] ));
			const int h0 = fabs( img->width * DotProduct( tv, basis[0] )) + fabs( img->height * DotProduct( tv, basis[1] ));
			const int x0 = DotProduct( p->position, tu ) + tu[3] - fa->texturemins[0] - w0 / 2;
			const int y0 = DotProduct( p->position, tv ) + tv[3] - fa->texturemins[1] - h0 / 2;
			o->tex_ofs = img->has_alpha ? te->alpha_ofs : te->ofs[0];
			o->texwidth = img->width;
			o->transparent = img->has_alpha ? 1 : 0;
			for( int mip = 0; mip < 4; mip++ )
			{
				rc_decalmip_t *dm = &o->mip[mip];
				const int surfwidth = fa->info->lightextents[0] >> mip, surfheight = fa->info->lightextents[1] >> mip;
				int s1 = 0, t1 = 0, s2 = img->width, t2 = img->height;
				int x = x0 >> mip, y = y0 >> mip, ww = w0 >> mip, hh = h0 >> mip;
				if( ww < 1 || hh < 1 )
					continue;
				if( x < 0 ) { s1 += ( -x ) * ( s2 - s1 ) / ww; x = 0; }
				if( x + ww > surfwidth ) { s2 -= ( x + ww - surfwidth ) * ( s2 - s1 ) / ww; ww = surfwidth - x; }
				if( y + hh > surfheight ) { t2 -= ( y + hh - surfheight ) * ( t2 - t1 ) / hh; hh = surfheight - y; }
				if( s1 < 0 ) s1 = 0;
				if( t1 < 0 ) t1 = 0;
				if( s2 > img->width ) s2 = img->width;
				if( t2 > img->height ) t2 = img->height;
				if( s1 >= s2 || t1 >= t2 || !ww )
					continue;
				int height = hh, skip = 0;
				if( y < 0 ) { skip = -y; height += y; y = 0; }
				if( ww < 0 || height <= 0 || hh == 0 )
					continue; /* the reference's loops would not run (or would run off the block) */
				unsigned fstep = ( s2 - s1 ) * 0x10000 / ww;
				if( ww == s2 - s1 )
					fstep = 0x10000;
				dm->valid = 1; dm->x = x; dm->y = y; dm->w = ww; dm->h = hh; dm->height = height; dm->skip = skip;
				dm->s1 = s1; dm->t1 = t1; dm->tspan = t2 - t1; dm->fstep = fstep;
			}
			n++; first[2 * i + 1]++;
		}
	}
	rc_decal_gen_ensure();
	RC_CHECK_DEV( rc_dev_set_decals( recs, n, first, rc_decal_gen, ns ));
	rc_free( recs ); rc_free( first );
}

static qboolean R_StudioFillAPI( struct engine_studio_api_s *api, struct r_studio_interface_s *pDefaultDraw ) { (void)api; (void)pDefaultDraw; return false; }
static void R_StudioSetDrawInterface( struct r_studio_interface_s *pDraw ) { (void)pDraw; }
static void R_SetSkyCloudsTextures( int solidskyTexture, int alphaskyTexture ) { (void)solidskyTexture; (void)alphaskyTexture; }
static void GL_SubdivideSurface( model_t *mod, msurface_t *fa ) { (void)mod; (void)fa; }
static qboolean Mod_ProcessRenderData( model_t *mod, qboolean create, const byte *buffer, size_t buffersize )
{
	(void)buffer; (void)buffersize;
	if( create && mod->type != mod_brush && mod->type != mod_studio && mod->type != mod_sprite && mod->type != mod_alias )
		return false;
	return true;
}
/* slot Mod_StudioLoadTextures (ref/soft/r_studio.c:3005-3115): every skin of a studio model goes through the engine's
 * image loader ("#model/skin.mdl" names tell it the buffer is an mstudiotexture_t whose pixels sit at the announced
 * pointer) into the texture table; mstudiotexture_t.index becomes the texture number the meshes bind */
static void Mod_StudioLoadTextures( model_t *mod, void *data )
{
	studiohdr_t *phdr = (studiohdr_t *)data;
	if( !phdr || phdr->textureindex <= 0 )
		return;
	mstudiotexture_t *ptexture = (mstudiotexture_t *)((byte *)phdr + phdr->textureindex );
	char mdlname[128], texname[300], *dot;
	rc_strncpy( mdlname, mod->name, sizeof( mdlname ));
	if(( dot = strrchr( mdlname, '.' )) != NULL ) *dot = 0;
	for( int i = 0; i < phdr->numtextures; i++ )
	{
		mstudiotexture_t *tx = &ptexture[i];
		int flags = 0;
		if( FBitSet( tx->flags, STUDIO_NF_NORMALMAP )) flags |= TF_NORMALMAP;
		if( FBitSet( tx->flags, STUDIO_NF_NOMIPS )) flags |= TF_NOMIPMAP;
		if( !strncasecmp( tx->name, "DM_Base", 7 ) || !strncasecmp( tx->name, "remap", 5 ))
			SetBits( tx->flags, STUDIO_NF_COLORMAP ); /* player colour remapping itself is not built */
		const char *base = strrchr( tx->name, '/' );
		char name[64];
		rc_strncpy( name, base ? base + 1 : tx->name, sizeof( name ));
		if(( dot = strrchr( name, '.' )) != NULL ) *dot = 0;
		gEngfuncs.Image_SetMDLPointer((byte *)phdr + tx->index );
		snprintf( texname, sizeof( texname ), "#%s/%s.mdl", mdlname, name );
		const size_t size = sizeof( mstudiotexture_t ) + tx->width * tx->height + 768;
		tx->index = GL_LoadTexture( texname, (byte *)tx, size, flags );
		if( !tx->index )
			tx->index = GL_FindTexture( REF_DEFAULT_TEXTURE );
	}
}
/* CL_DrawParticles (ref/soft/r_part.c:47): every live particle becomes a draw item; its quad is built per view on the device
 * (rc_particle_corners).  Colour from the engine's palette, faded over the last 1/16 s of the particle's life. */
static void CL_DrawParticles( double frametime, particle_t *particles, float partsize )
{
	if( !particles )
		return;
	const color24 *palette = (const color24 *)gEngfuncs.EngineGetParm( PARM_GET_PALETTE_PTR, 0 );
	GL_SetRenderMode( kRenderTransAdd );
	GL_Bind( XASH_TEXTURE0, rc.particleTexture );
	for( particle_t *p = particles; p; p = p->next )
	{
		if(( p->type != pt_blob ) || ( p->unused == 255 ))
		{
			p->color = bound( 0, p->color, 255 );
			const color24 color = palette ? palette[p->color] : (color24){ 255, 255, 255 };
			int alpha = 255 * ( p->die - gp_cl->time ) * 16.0f;
			if( alpha > 255 || p->type == pt_static )
				alpha = 255;
			rc_tri_color4f( 1.0f * alpha / 255 / 255 * color.r, 1.0f * alpha / 255 / 255 * color.g, 1.0f * alpha / 255 / 255 * color.b, 1.0f );
			rc_item_t *it = rc_new_item( RC_ITEM_PARTICLE );
			it->count = 2;
			it->light = (int32_t)( rc.tri.light << 8 );
			it->scale = partsize;
			VectorCopy( p->org, it->origin );
			TriTexCoord2f( 0.0f, 0.0f ); it->s0 = rc.tri.s; it->t0 = rc.tri.t;
			TriTexCoord2f( 1.0f, 1.0f ); it->s1 = rc.tri.s; it->t1 = rc.tri.t;
		}
		gEngfuncs.CL_ThinkParticle( frametime, p );
	}
}

/*
 * CL_DrawTracers (ref/soft/r_part.c:137): a tracer is a streak from p->org along p->vel; its quad is as wide as
 * gTracerSize says, at right angles to the streak's direction ON THE SCREEN of the view that draws it (R_WorldToScreen of
 * both ends).  The host records one item per tracer - start, streak vector, half width, colour - and every view builds
 * its own quad on the device (rc_tracer_corners); ref_soft culls none (CL_CullTracer returns false).  The movement the
 * function applies to the client's list afterwards stays here.
 */
static const float rc_tracer_size[11] = { 1.5f, 0.5f, 1.0f, 1.0f, 1.0f, 1.0f, 1.0f, 1.0f, 1.0f, 1.0f, 1.0f };
static color24 rc_tracer_colors[12] =
{
	{ 255, 255, 255 }, { 255, 0, 0 }, { 0, 255, 0 }, { 0, 0, 255 },
	{ 0, 0, 0 },       /* TRACER_COLORINDEX_DEFAULT: from tracerred / tracergreen / tracerblue / traceralpha */
	{ 255, 167, 17 }, { 255, 130, 90 }, { 55, 60, 144 }, { 255, 130, 90 }, { 255, 140, 90 }, { 200, 130, 90 }, { 255, 120, 70 },
};

static void CL_DrawTracers( double frametime, particle_t *tracers )
{
	if( FBitSet( tracerred->flags | tracergreen->flags | tracerblue->flags | traceralpha->flags, FCVAR_CHANGED ))
	{
		color24 *custom = &rc_tracer_colors[4];
		custom->r = (byte)( tracerred->value * traceralpha->value * 255 );
		custom->g = (byte)( tracergreen->value * traceralpha->value * 255 );
		custom->b = (byte)( tracerblue->value * traceralpha->value * 255 );
		ClearBits( tracerred->flags, FCVAR_CHANGED );
		ClearBits( tracergreen->flags, FCVAR_CHANGED );
		ClearBits( tracerblue->flags, FCVAR_CHANGED );
		ClearBits( traceralpha->flags, FCVAR_CHANGED );
	}
	if( !tracers )
		return;
	GL_SetRenderMode( kRenderTransAdd );
	/* TriSpriteTexture( GetDefaultSprite( REF_DOT_SPRITE ), 0 ), r_triapi.c:282 */
	model_t *dot = gEngfuncs.GetDefaultSprite( REF_DOT_SPRITE );
	if( !dot || dot->type != mod_sprite || !dot->cache.data )
		return;
	int texnum = gEngfuncs.R_GetSpriteFrame( dot, 0, 0.0f )->gl_texturenum;
	if( texnum == 0 )
		return;
	if( texnum < 0 || texnum >= RC_MAX_TEXTURES )
		texnum = 0;
	GL_Bind( XASH_TEXTURE0, texnum );

	const movevars_t *mv = (const movevars_t *)gEngfuncs.EngineGetParm( PARM_GET_MOVEVARS_PTR, 0 );
	const float gravity = frametime * ( mv ? mv->gravity : 800.0f );
	float scale = 1.0 - ( frametime * 0.9 );
	if( scale < 0.0f )
		scale = 0.0f;
	for( particle_t *p = tracers; p; p = p->next )
	{
		float atten = ( p->die - gp_cl->time );
		if( atten > 0.1f )
			atten = 0.1f;
		vec3_t delta;
		VectorScale( p->vel, ( p->ramp * atten ), delta );
		const short alpha = p->unused;
		if( p->color < 0 || p->color >= (int)( sizeof( rc_tracer_colors ) / sizeof( rc_tracer_colors[0] )))
			p->color = TRACER_COLORINDEX_DEFAULT;
		const color24 color = rc_tracer_colors[p->color];
		rc_tri_color4f( 1.0f * alpha / 255 / 255 * color.r, 1.0f * alpha / 255 / 255 * color.g, 1.0f * alpha / 255 / 255 * color.b, 1.0f );
		rc_item_t *it = rc_new_item( RC_ITEM_TRACER );
		rc.ntracers++;
		it->count = 2;
		it->light = (int32_t)( rc.tri.light << 8 );
		it->scale = ( (unsigned)p->type < sizeof( rc_tracer_size ) / sizeof( rc_tracer_size[0] )) ? rc_tracer_size[p->type] : 1.0f; /* (the reference indexes past its table for the custom particle types) */
		VectorCopy( p->org, it->origin );
		VectorCopy( delta, it->angles );
		TriTexCoord2f( 0.0f, 0.0f ); it->s0 = rc.tri.s; it->t0 = rc.tri.t;
		TriTexCoord2f( 1.0f, 0.8f ); it->s1 = rc.tri.s; it->t1 = rc.tri.t;

		/* the client's list moves on */
		VectorMA( p->org, frametime, p->vel, p->org );
		if( p->type == pt_grav )
		{
			p->vel[0] *= scale;
			p->vel[1] *= scale;
			p->vel[2] -= gravity;
			p->unused = 255 * ( p->die - gp_cl->time ) * 2;
			if( p->unused > 255 )
				p->unused = 255;
		}
		else if( p->type == pt_slowgrav )
		{
			p->vel[2] = gravity * 0.05;
		}
	}
}
static void CL_DrawBeams( int fTrans, BEAM *beams ) { (void)fTrans; (void)beams; }

static intptr_t RefGetParm( int parm, int arg )
{
	rc_image_t *t = ( arg >= 0 && arg < RC_MAX_TEXTURES ) ? &rc.images[arg] : &rc.images[0];
	switch( parm )
	{
	case PARM_TEX_WIDTH: return t->width;
	case PARM_TEX_HEIGHT: return t->height;
	case PARM_TEX_SRC_WIDTH: return t->srcWidth;
	case PARM_TEX_SRC_HEIGHT: return t->srcHeight;
	case PARM_TEX_MIPCOUNT: return RC_IMAGE_LIVE( t ) ? 4 : 0;
	case PARM_TEX_DEPTH: return 1;
	case PARM_TEX_FLAGS: return t->flags;
	case PARM_LIGHTSTYLEVALUE: return rc.lightstylevalue[bound( 0, arg, MAX_LIGHTSTYLES - 1 )];
	case PARM_GL_CONTEXT_TYPE: return CONTEXT_TYPE_SOFTWARE;
	case PARM_NUMENTITIES: return 0;
	default: return 0;
	}
}

static void R_GetDetailScaleForTexture( int texture, float *xScale, float *yScale ) { (void)texture; if( xScale ) *xScale = 1.0f; if( yScale ) *yScale = 1.0f; }
static void R_SetDetailScaleForTexture( int texture, float xScale, float yScale ) { (void)texture; (void)xScale; (void)yScale; }
static void R_OverrideTextureSourceSize( unsigned int texnum, unsigned int srcWidth, unsigned int srcHeight )
{
	if( texnum < RC_MAX_TEXTURES ) { rc.images[texnum].srcWidth = srcWidth; rc.images[texnum].srcHeight = srcHeight; }
}
static void GL_UpdateTexture( int texnum, int cols, int rows, int width, int height, const byte *buffer, pixformat_t fmt )
{ (void)texnum; (void)cols; (void)rows; (void)width; (void)height; (void)buffer; (void)fmt; }
static void GL_OrthoBounds( const float *mins, const float *maxs ) { (void)mins; (void)maxs; }

static qboolean R_SpeedsMessage( char *out, size_t size )
{
	uint64_t c[RC_CTR_COUNT];
	if( !r_speeds->value || rc_dev_read_counters( c, RC_CTR_COUNT ) != 0 )
		return false;
	snprintf( out, size, "ref_cuda: %llu views %llu faces %llu edges %llu spans, %llu cache blocks built\n",
		(unsigned long long)c[RC_CTR_VIEWS], (unsigned long long)c[RC_CTR_SURFS], (unsigned long long)c[RC_CTR_EDGES],
		(unsigned long long)c[RC_CTR_SPANS], (unsigned long long)c[RC_CTR_CACHE_BLOCKS] );
	return true;
}
static byte *Mod_GetCurrentVis( void ) { return NULL; }
static void R_FillRenderAPI( struct render_api_s *api ) { (void)api; }
static void R_FillTriAPI( struct triangleapi_s *api ) { (void)api; }
static void VGUI_SetupDrawing( qboolean rect ) { (void)rect; }

/* designated initialisers: the slot order is the ABI (engine/ref_api.h:527-658), names keep it checkable */
static const ref_interface_t gReffuncs =
{
	.R_Init = R_Init, .R_Shutdown = R_Shutdown, .R_GetConfigName = R_GetConfigName, .R_SetDisplayTransform = R_SetDisplayTransform,
	.GL_SetupAttributes = GL_SetupAttributes, .GL_InitExtensions = GL_InitExtensions, .GL_ClearExtensions = GL_ClearExtensions,
	.R_GammaChanged = R_GammaChanged, .R_BeginFrame = R_BeginFrame, .R_RenderScene = R_RenderScene, .R_EndFrame = R_EndFrame,
	.R_PushScene = R_PushScene, .R_PopScene = R_PopScene, .GL_BackendStartFrame = GL_BackendStartFrame, .GL_BackendEndFrame = GL_BackendEndFrame,
	.R_ClearScreen = R_ClearScreen, .R_AllowFog = R_AllowFog, .GL_SetRenderMode = GL_SetRenderMode,
	.R_AddEntity = R_AddEntity, .R_ProcessEntData = R_ProcessEntData, .R_ShowTextures = R_ShowTextures,
	.R_GetTextureOriginalBuffer = R_GetTextureOriginalBuffer, .GL_LoadTextureFromBuffer = GL_LoadTextureFromBuffer,
	.GL_ProcessTexture = GL_ProcessTexture, .R_SetupSky = R_SetupSky,
	.R_Set2DMode = R_Set2DMode, .R_DrawStretchPic = R_DrawStretchPic, .FillRGBA = FillRGBA, .WorldToScreen = WorldToScreen,
	.VID_ScreenShot = VID_ScreenShot, .VID_CubemapShot = VID_CubemapShot, .R_LightPoint = R_LightPoint,
	.R_DecalShoot = R_DecalShoot, .R_DecalRemoveAll = R_DecalRemoveAll, .R_CreateDecalList = R_CreateDecalList, .R_ClearAllDecals = R_ClearAllDecals,
	.R_StudioEstimateFrame = R_StudioEstimateFrame, .R_StudioLerpMovement = R_StudioLerpMovement,
	.R_StudioFillAPI = R_StudioFillAPI, .R_StudioSetDrawInterface = R_StudioSetDrawInterface,
	.R_SetSkyCloudsTextures = R_SetSkyCloudsTextures, .GL_SubdivideSurface = GL_SubdivideSurface, .CL_RunLightStyles = CL_RunLightStyles,
	.Mod_ProcessRenderData = Mod_ProcessRenderData, .Mod_StudioLoadTextures = Mod_StudioLoadTextures,
	.CL_DrawParticles = CL_DrawParticles, .CL_DrawTracers = CL_DrawTracers, .CL_DrawBeams = CL_DrawBeams,
	.RefGetParm = RefGetParm,
	.R_GetDetailScaleForTexture = R_GetDetailScaleForTexture, .R_SetDetailScaleForTexture = R_SetDetailScaleForTexture,
	.GL_CreateTexture = GL_CreateTexture, .GL_FindTexture = GL_FindTexture, .GL_TextureName = GL_TextureName, .GL_TextureData = GL_TextureData,
	.GL_LoadTexture = GL_LoadTexture, .GL_FreeTexture = GL_FreeTexture, .R_OverrideTextureSourceSize = R_OverrideTextureSourceSize,
	.GL_UpdateTexture = GL_UpdateTexture, .GL_Bind = GL_Bind,
	.GL_RenderFrame = R_RenderFrame, .GL_OrthoBounds = GL_OrthoBounds, .R_SpeedsMessage = R_SpeedsMessage,
	.Mod_GetCurrentVis = Mod_GetCurrentVis, .R_NewMap = R_NewMap, .R_ClearScene = R_ClearScene,
	.TriRenderMode = TriRenderMode, .Begin = TriBegin, .End = TriEnd, .Color4f = rc_tri_color4f, .Color4ub = rc_tri_color4ub, /* _TriColor4f / _TriColor4ub (r_context.c:523) */
	.Vertex3fv = TriVertex3fv, .Vertex3f = TriVertex3f, .CullFace = TriCullFace,
	.R_FillRenderAPI = R_FillRenderAPI, .R_FillTriAPI = R_FillTriAPI, .VGUI_SetupDrawing = VGUI_SetupDrawing,
};

/* engine/ref_api.h:660; handshake as ref/common/ref_context.c:57-81 */
int GetRefAPI( int version, ref_interface_t *funcs, ref_api_t *engfuncs, ref_globals_t *globals )
{
	if( version != REF_API_VERSION )
		return 0;

	*funcs = gReffuncs;
	gEngfuncs = *engfuncs;
	gpGlobals = globals;

	gp_cl = (ref_client_t *)gEngfuncs.EngineGetParm( PARM_GET_CLIENT_PTR, 0 );
	gp_host = (ref_host_t *)gEngfuncs.EngineGetParm( PARM_GET_HOST_PTR, 0 );

	RETRIEVE_ENGINE_SHARED_CVAR_LIST();

	return REF_API_VERSION;
}
