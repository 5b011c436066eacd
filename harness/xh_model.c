/*
 * xh_model.c - BSP30 -> model_t, PVS services, .mip decoding, gamma tables.
 *
 * These are the engine-side inputs of the renderer hot path:
 * each function names the engine routine whose observable result it restates.
 * They are shared by every renderer under this host, so they cannot bias a
 * parity comparison; the gamma tables additionally carry the reference's own
 * known-answer test (engine/client/gamma.c:245-268) in tests/.
 */
#define _GNU_SOURCE
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "xh_internal.h"
#include "studio.h"
#include "wadfile.h"

/* ------------------------------------------------------------------ */
/* gamma tables - engine/client/gamma.c:35-80 BuildGammaTable            */
/* cast placement matters (gamma.c:54,72): float d from double pow, etc. */

void xh_build_gamma( float gamma, float brightness, float texgamma, float lightgamma )
{
	float g1 = ( gamma != 0.0 ) ? (float)( 1.0 / gamma ) : (float)0.4;
	float g2 = g1 * texgamma;
	float g3;

	if( brightness <= 0.0 ) g3 = 0.125;
	else if( brightness <= 1.0 ) g3 = 0.125 - brightness * brightness * 0.075;
	else g3 = 0.05;

	for( int i = 0; i < 256; i++ )
	{
		float d = pow( i / 255.0, (double)g2 );
		int inf = d * 255.0f;
		xh.texgamma[i] = inf < 0 ? 0 : ( inf > 255 ? 255 : inf );
	}

	for( int i = 0; i < 1024; i++ )
	{
		float f = pow( i / 1023.0, (double)lightgamma );

		if( brightness > 1.0 ) f *= brightness;
		if( f <= g3 ) f = ( f / g3 ) * 0.125;
		else f = (( f - g3 ) / ( 1.0 - g3 )) * 0.875 + 0.125;

		double d = pow((double)f, (double)g1 );
		int inf = d * 1023.0;
		xh.lightgamma[i] = inf < 0 ? 0 : ( inf > 1023 ? 1023 : inf );
		xh.lineargamma[i] = pow( i / 1023.0, (double)gamma ) * 1023.0;
		xh.screengamma[i] = pow( i / 1023.0, 1.0 / gamma ) * 1023.0;
	}
}

const void *xh_gamma_table( int which, int *count )
{
	switch( which )
	{
	case 0: if( count ) *count = 256; return xh.texgamma;
	case 1: if( count ) *count = 1024; return xh.lightgamma;
	case 2: if( count ) *count = 1024; return xh.lineargamma;
	case 3: if( count ) *count = 1024; return xh.screengamma;
	}
	return NULL;
}

/* ------------------------------------------------------------------ */
/* images: FS_LoadImage for "#map:name.mip" buffers (engine/common/imagelib/img_wad.c Image_LoadMIP) */
/* 8-bit indexed miptex + 256-entry palette after the 4th mip -> RGBA8    */

extern byte *xh_mdl_pointer;

/* engine/common/imagelib/img_wad.c Image_LoadMDL: a studio model skin - mstudiotexture_t header, 8-bit indices at the
 * pointer the renderer announced through Image_SetMDLPointer, the palette right after them */
static rgbdata_t *xh_load_mdl_skin( const byte *buffer )
{
	const mstudiotexture_t *tx = (const mstudiotexture_t *)buffer;
	const byte *src = xh_mdl_pointer;
	if( !src || tx->width <= 0 || tx->height <= 0 || tx->width > 4096 || tx->height > 4096 ) return NULL;
	const size_t pixels = (size_t)tx->width * tx->height;
	const byte *pal = src + pixels;
	rgbdata_t *pic = calloc( 1, sizeof( *pic ));
	pic->width = tx->width; pic->height = tx->height; pic->depth = 1;
	pic->type = PF_RGBA_32; pic->flags = IMAGE_HAS_COLOR; pic->size = pixels * 4; pic->numMips = 1;
	pic->buffer = malloc( pic->size );
	const qboolean masked = FBitSet( tx->flags, STUDIO_NF_MASKED );
	for( size_t i = 0; i < pixels; i++ )
	{
		const byte idx = src[i];
		byte *out = pic->buffer + i * 4;
		if( masked && idx == 255 ) { out[0] = out[1] = out[2] = out[3] = 0; pic->flags |= IMAGE_HAS_ALPHA | IMAGE_ONEBIT_ALPHA; }
		else { out[0] = pal[idx * 3]; out[1] = pal[idx * 3 + 1]; out[2] = pal[idx * 3 + 2]; out[3] = 255; }
	}
	return pic;
}

rgbdata_t *xh_FS_LoadImage( const char *filename, const byte *buffer, size_t size )
{
	xh.callbacks++;
	const size_t fl = filename ? strlen( filename ) : 0;
	if( buffer && fl > 4 && !strcmp( filename + fl - 4, ".mdl" ))
		return xh_load_mdl_skin( buffer );
	if( !buffer || size < sizeof( mip_t )) return NULL;

	const mip_t *mt = (const mip_t *)buffer;
	uint w = mt->width, h = mt->height;
	if( !w || !h || w > 4096 || h > 4096 || !mt->offsets[0] ) return NULL;

	size_t pixels = (size_t)w * h;
	size_t palofs = mt->offsets[0] + (( pixels * 85 ) >> 6 ) + sizeof( short );
	if( palofs + 768 > size ) return NULL;

	const byte *src = buffer + mt->offsets[0];
	const byte *pal = buffer + palofs;
	const char *base = strrchr( filename, ':' );
	base = base ? base + 1 : filename;
	qboolean transparent = base[0] == '{';

	rgbdata_t *pic = calloc( 1, sizeof( *pic ));
	pic->width = w; pic->height = h; pic->depth = 1;
	pic->type = PF_RGBA_32;
	pic->flags = IMAGE_HAS_COLOR;
	pic->size = pixels * 4;
	pic->buffer = malloc( pic->size );
	pic->numMips = 1;

	for( size_t i = 0; i < pixels; i++ )
	{
		byte idx = src[i];
		byte *out = pic->buffer + i * 4;
		if( transparent && idx == 255 )
		{
			out[0] = out[1] = out[2] = out[3] = 0;
			pic->flags |= IMAGE_HAS_ALPHA | IMAGE_ONEBIT_ALPHA;
		}
		else
		{
			out[0] = pal[idx * 3 + 0]; out[1] = pal[idx * 3 + 1]; out[2] = pal[idx * 3 + 2]; out[3] = 255;
		}
	}
	return pic;
}

rgbdata_t *xh_FS_CopyImage( const rgbdata_t *in )
{
	if( !in ) return NULL;
	rgbdata_t *out = malloc( sizeof( *out ));
	*out = *in;
	out->palette = NULL;
	if( in->buffer ) { out->buffer = malloc( in->size ); memcpy( out->buffer, in->buffer, in->size ); }
	return out;
}

void xh_FS_FreeImage( rgbdata_t *pic )
{
	if( !pic ) return;
	free( pic->buffer );
	free( pic );
}

/* ------------------------------------------------------------------ */
/* PVS services                                                          */

/* engine/common/mod_bmodel.c:1426 Mod_SampleSizeForFace */
int xh_Mod_SampleSizeForFace( const msurface_t *surf )
{
	if( !surf || !surf->texinfo ) return LM_SAMPLE_SIZE;
	if( FBitSet( surf->texinfo->flags, TEX_WORLD_LUXELS )) return 1;
	if( FBitSet( surf->texinfo->flags, TEX_EXTRA_LIGHTMAP )) return LM_SAMPLE_EXTRASIZE;
	if( surf->texinfo->faceinfo ) return surf->texinfo->faceinfo->texture_step;
	return LM_SAMPLE_SIZE;
}

/* engine/common/mod_bmodel.c:1116 Mod_PointInLeaf: "<= 0" goes to the back child */
mleaf_t *xh_Mod_PointInLeaf( const vec3_t p, mnode_t *node, model_t *mod )
{
	xh.callbacks++;
	while( node->contents >= 0 )
		node = node_child( node, PlaneDiff( p, node->plane ) <= 0, mod );
	return (mleaf_t *)node;
}

/* engine/common/mod_bmodel.c:1054 Mod_DecompressPVS: zero-run RLE, NULL row == all visible */
static void xh_decompress_vis( byte *out, const byte *in, size_t visbytes )
{
	byte *dst = out, *end = out + visbytes;
	if( !in ) { memset( out, 0xFF, visbytes ); return; }
	while( dst < end )
	{
		if( *in ) { *dst++ = *in++; continue; }
		size_t run = in[1];
		if( run > (size_t)( end - dst )) run = end - dst;
		memset( dst, 0, run );
		dst += run; in += 2;
	}
}

/* engine/common/mod_bmodel.c:1160 Mod_FatPVS_RecursiveBSPNode */
static void xh_fatpvs_node( const float *org, float radius, byte *vis, int bytes, mnode_t *node )
{
	while( node->contents >= 0 )
	{
		float d = PlaneDiff( org, node->plane );
		if( d > radius ) node = node_child( node, 0, xh.world );
		else if( d < -radius ) node = node_child( node, 1, xh.world );
		else
		{
			xh_fatpvs_node( org, radius, vis, bytes, node_child( node, 0, xh.world ));
			node = node_child( node, 1, xh.world );
		}
	}
	mleaf_t *leaf = (mleaf_t *)node;
	if( leaf->cluster >= 0 )
	{
		xh_decompress_vis( xh.visscratch, leaf->compressed_vis, xh.visbytes );
		for( int i = 0; i < bytes; i++ ) vis[i] |= xh.visscratch[i];
	}
}

/* engine/client/cl_render.c:21 R_FatPVS -> mod_bmodel.c:1203 Mod_FatPVS( ..., world.visbytes, merge, fullvis, false ) */
int xh_R_FatPVS( const float *org, float radius, byte *visbuffer, qboolean merge, qboolean fullvis )
{
	xh.callbacks++;
	if( !xh.world ) return 0;
	mleaf_t *leaf = xh_Mod_PointInLeaf( org, xh.world->nodes, xh.world );
	int bytes = (int)xh.visbytes;

	if( fullvis || !xh.world->visdata || !leaf || leaf->cluster < 0 )
	{
		memset( visbuffer, 0xFF, bytes );
		return bytes;
	}
	if( !merge ) memset( visbuffer, 0, bytes );
	xh_fatpvs_node( org, radius, visbuffer, bytes, xh.world->nodes );
	return bytes;
}

qboolean xh_Mod_BoxVisible( const vec3_t mins, const vec3_t maxs, const byte *visbits )
{
	(void)mins; (void)maxs; (void)visbits;
	return true; /* conservative; only used for entity culling */
}

int xh_point_in_leaf( const float *p, int *contents, int *cluster )
{
	if( !xh.world ) return -1;
	mleaf_t *leaf = xh_Mod_PointInLeaf( p, xh.world->nodes, xh.world );
	if( contents ) *contents = leaf->contents;
	if( cluster ) *cluster = leaf->cluster;
	return (int)( leaf - xh.world->leafs );
}

/* ------------------------------------------------------------------ */
/* BSP30 loader (engine/common/mod_bmodel.c:2424-3970,4145-4288)          */

typedef struct { const byte *base; size_t len; const dheader_t *hdr; } xh_bsp_t;

static const void *xh_lump( const xh_bsp_t *b, int lump, size_t elem, int *count )
{
	const dlump_t *l = &b->hdr->lumps[lump];
	if( l->fileofs < 0 || l->filelen < 0 || (size_t)l->fileofs + l->filelen > b->len )
		return NULL;
	*count = (int)( l->filelen / elem );
	return b->base + l->fileofs;
}

/* mod_bmodel.c:2579 Mod_LooksLikeWaterTexture (Half-Life rules) */
static qboolean xh_is_water_name( const char *n )
{
	if(( n[0] == '*' && strcasecmp( n, REF_DEFAULT_TEXTURE )) || n[0] == '!' ) return true;
	return !strncmp( n, "water", 5 ) || !strncasecmp( n, "laser", 5 );
}

/* mod_bmodel.c:2967-3100 Mod_SequenceAnimatedTexture: "+0name".."+9name" / "+aname".. chains */
static void xh_sequence_textures( model_t *mod )
{
	for( int i = 0; i < mod->numtextures; i++ )
	{
		texture_t *tx = mod->textures[i];
		if( !tx || ( tx->name[0] != '-' && tx->name[0] != '+' ) || tx->anim_next ) continue;

		texture_t *anims[10] = { 0 }, *alts[10] = { 0 };
		int max = 0, altmax = 0;
		for( int j = 0; j < mod->numtextures; j++ )
		{
			texture_t *t2 = mod->textures[j];
			if( !t2 || t2->name[0] != tx->name[0] || strcmp( t2->name + 2, tx->name + 2 )) continue;
			int c = t2->name[1];
			if( c >= 'a' && c <= 'z' ) c -= 'a' - 'A';
			if( c >= '0' && c <= '9' ) { c -= '0'; anims[c] = t2; if( c + 1 > max ) max = c + 1; }
			else if( c >= 'A' && c <= 'J' ) { c -= 'A'; alts[c] = t2; if( c + 1 > altmax ) altmax = c + 1; }
		}
		for( int j = 0; j < max; j++ )
		{
			texture_t *t2 = anims[j];
			if( !t2 ) return; /* broken chain: leave unsequenced, renderer falls back to base */
			t2->anim_total = max * 2; /* ANIM_CYCLE */
			t2->anim_min = j * 2; t2->anim_max = ( j + 1 ) * 2;
			t2->anim_next = anims[( j + 1 ) % max];
			if( altmax ) t2->alternate_anims = alts[0];
		}
		for( int j = 0; j < altmax; j++ )
		{
			texture_t *t2 = alts[j];
			if( !t2 ) return;
			t2->anim_total = altmax * 2;
			t2->anim_min = j * 2; t2->anim_max = ( j + 1 ) * 2;
			t2->anim_next = alts[( j + 1 ) % altmax];
			if( max ) t2->alternate_anims = anims[0];
		}
	}
}

static int xh_load_textures( model_t *mod, const xh_bsp_t *b, const char *mapname )
{
	const dlump_t *l = &b->hdr->lumps[LUMP_TEXTURES];
	if( l->filelen < 4 ) { mod->numtextures = 0; return 0; }
	const byte *lump = b->base + l->fileofs;
	const dmiptexlump_t *in = (const dmiptexlump_t *)lump;

	mod->numtextures = in->nummiptex;
	mod->textures = xh_alloc( sizeof( texture_t * ) * in->nummiptex );

	for( int i = 0; i < in->nummiptex; i++ )
	{
		texture_t *tx = xh_alloc( sizeof( *tx ));
		mod->textures[i] = tx;
		int ofs = in->dataofs[i];
		if( ofs < 0 || (size_t)ofs + sizeof( mip_t ) > (size_t)l->filelen )
		{
			strcpy( tx->name, "*default" ); tx->width = tx->height = 16;
			tx->gl_texturenum = xh.ref.GL_FindTexture( REF_DEFAULT_TEXTURE );
			continue;
		}
		const mip_t *mt = (const mip_t *)( lump + ofs );
		for( int k = 0; k < 15 && mt->name[k]; k++ )
			tx->name[k] = ( mt->name[k] >= 'A' && mt->name[k] <= 'Z' ) ? mt->name[k] + 32 : mt->name[k];
		tx->width = mt->width; tx->height = mt->height;

		uint flags = 0;
		if( xh_is_water_name( tx->name )) flags |= TF_KEEP_SOURCE | TF_EXPAND_SOURCE;
		size_t pixels = (size_t)mt->width * mt->height;
		size_t size = sizeof( mip_t ) + (( pixels * 85 ) >> 6 ) + sizeof( short ) + 768;
		if((size_t)ofs + size > (size_t)l->filelen ) size = l->filelen - ofs;

		char texname[128];
		snprintf( texname, sizeof( texname ), "#%s:%s.mip", mapname, tx->name );
		tx->gl_texturenum = xh.ref.GL_LoadTexture( texname, (const byte *)mt, size, flags );
		if( !tx->gl_texturenum )
			tx->gl_texturenum = xh.ref.GL_FindTexture( REF_DEFAULT_TEXTURE );
	}
	xh_sequence_textures( mod );
	return 0;
}

/* mod_bmodel.c:1590 Mod_CalcSurfaceExtents (+ :1543 Mod_LightMatrixFromTexMatrix for the plain case) */
static void xh_calc_extents( model_t *mod, msurface_t *surf )
{
	float mins[2] = { 999999, 999999 }, maxs[2] = { -999999, -999999 };
	mextrasurf_t *info = surf->info;
	int sample_size = xh_Mod_SampleSizeForFace( surf );
	mtexinfo_t *tex = surf->texinfo;

	for( int i = 0; i < 2; i++ )
		for( int j = 0; j < 4; j++ )
			info->lmvecs[i][j] = tex->vecs[i][j];

	for( int i = 0; i < surf->numedges; i++ )
	{
		int e = mod->surfedges[surf->firstedge + i];
		mvertex_t *v = e >= 0 ? &mod->vertexes[mod->edges16[e].v[0]] : &mod->vertexes[mod->edges16[-e].v[1]];
		for( int j = 0; j < 2; j++ )
		{
			float val = DotProductPrecise( v->position, tex->vecs[j] ) + tex->vecs[j][3];
			if( val < mins[j] ) mins[j] = val;
			if( val > maxs[j] ) maxs[j] = val;
		}
	}

	for( int i = 0; i < 2; i++ )
	{
		int bmins = floor( mins[i] / sample_size );
		int bmaxs = ceil( maxs[i] / sample_size );
		surf->texturemins[i] = bmins * sample_size;
		surf->extents[i] = ( bmaxs - bmins ) * sample_size;
		info->lightmapmins[i] = surf->texturemins[i];
		info->lightextents[i] = surf->extents[i];
	}
}

static void xh_calc_bounds( model_t *mod, msurface_t *surf )
{
	mextrasurf_t *info = surf->info;
	info->mins[0] = info->mins[1] = info->mins[2] = 999999;
	info->maxs[0] = info->maxs[1] = info->maxs[2] = -999999;
	for( int i = 0; i < surf->numedges; i++ )
	{
		int e = mod->surfedges[surf->firstedge + i];
		mvertex_t *v = e >= 0 ? &mod->vertexes[mod->edges16[e].v[0]] : &mod->vertexes[mod->edges16[-e].v[1]];
		for( int j = 0; j < 3; j++ )
		{
			if( v->position[j] < info->mins[j] ) info->mins[j] = v->position[j];
			if( v->position[j] > info->maxs[j] ) info->maxs[j] = v->position[j];
		}
	}
	for( int j = 0; j < 3; j++ ) info->origin[j] = ( info->mins[j] + info->maxs[j] ) * 0.5f;
}

static void xh_set_parent( model_t *mod, mnode_t *node, mnode_t *parent )
{
	node->parent = parent;
	if( node->contents < 0 ) return;
	xh_set_parent( mod, node_child( node, 0, mod ), node );
	xh_set_parent( mod, node_child( node, 1, mod ), node );
}

int xh_load_map( const void *data, size_t len, const char *name )
{
	xh_bsp_t b = { data, len, data };
	int n;

	if( len < sizeof( dheader_t ) || b.hdr->version != HLBSP_VERSION ) return -1;

	model_t *mod = xh_alloc( sizeof( *mod ));
	snprintf( mod->name, sizeof( mod->name ), "maps/%s.bsp", name );
	mod->type = mod_brush;
	mod->mempool = 1;
	xh.world = mod;

	/* submodels first: leaf loading needs visleafs (Mod_LoadBmodelLumps order, mod_bmodel.c:4230-4260) */
	const dmodel_t *dm = xh_lump( &b, LUMP_MODELS, sizeof( dmodel_t ), &n );
	if( !dm || n < 1 ) return -2;
	mod->numsubmodels = n;
	mod->submodels = xh_alloc( sizeof( dmodel_t ) * n );
	memcpy( mod->submodels, dm, sizeof( dmodel_t ) * n );

	/* entities */
	{
		const dlump_t *l = &b.hdr->lumps[LUMP_ENTITIES];
		mod->entities = xh_alloc( l->filelen + 1 );
		memcpy( mod->entities, b.base + l->fileofs, l->filelen );
	}

	/* planes (:2424) */
	const dplane_t *dp = xh_lump( &b, LUMP_PLANES, sizeof( dplane_t ), &n );
	mod->numplanes = n;
	mod->planes = xh_alloc( sizeof( mplane_t ) * n );
	for( int i = 0; i < n; i++ )
	{
		mplane_t *out = &mod->planes[i];
		for( int j = 0; j < 3; j++ )
		{
			out->normal[j] = dp[i].normal[j];
			if( out->normal[j] < 0.0f ) out->signbits |= 1 << j;
		}
		out->dist = dp[i].dist;
		out->type = dp[i].type;
	}

	/* vertexes, edges, surfedges */
	const dvertex_t *dv = xh_lump( &b, LUMP_VERTEXES, sizeof( dvertex_t ), &n );
	mod->numvertexes = n;
	mod->vertexes = xh_alloc( sizeof( mvertex_t ) * n );
	memcpy( mod->vertexes, dv, sizeof( mvertex_t ) * n );

	const dedge_t *de = xh_lump( &b, LUMP_EDGES, sizeof( dedge_t ), &n );
	mod->numedges = n;
	mod->edges16 = xh_alloc( sizeof( medge16_t ) * n );
	for( int i = 0; i < n; i++ ) { mod->edges16[i].v[0] = de[i].v[0]; mod->edges16[i].v[1] = de[i].v[1]; }

	const dsurfedge_t *ds = xh_lump( &b, LUMP_SURFEDGES, sizeof( dsurfedge_t ), &n );
	mod->numsurfedges = n;
	mod->surfedges = xh_alloc( sizeof( int ) * n );
	memcpy( mod->surfedges, ds, sizeof( int ) * n );

	/* textures need cl.models[1] for nothing, but the renderer's table must exist: R_Init already ran */
	xh_load_textures( mod, &b, name );

	/* texinfo (:3247) */
	const dtexinfo_t *dt = xh_lump( &b, LUMP_TEXINFO, sizeof( dtexinfo_t ), &n );
	mod->numtexinfo = n;
	mod->texinfo = xh_alloc( sizeof( mtexinfo_t ) * n );
	for( int i = 0; i < n; i++ )
	{
		memcpy( mod->texinfo[i].vecs, dt[i].vecs, sizeof( dt[i].vecs ));
		int miptex = dt[i].miptex;
		if( miptex < 0 || miptex >= mod->numtextures ) miptex = 0;
		mod->texinfo[i].texture = mod->textures[miptex];
		mod->texinfo[i].flags = dt[i].flags;
	}

	/* lighting: BSP30 stores RGB triplets, lightofs is in bytes (:3883-3968) */
	{
		const dlump_t *l = &b.hdr->lumps[LUMP_LIGHTING];
		if( l->filelen > 0 )
		{
			mod->lightdata = xh_alloc( l->filelen );
			memcpy( mod->lightdata, b.base + l->fileofs, l->filelen );
			mod->flags |= MODEL_COLORED_LIGHTING;
		}
	}

	/* surfaces (:3291) */
	const dface_t *df = xh_lump( &b, LUMP_FACES, sizeof( dface_t ), &n );
	mod->numsurfaces = n;
	mod->surfaces = xh_alloc( sizeof( msurface_t ) * n );
	mextrasurf_t *info = xh_alloc( sizeof( mextrasurf_t ) * n );
	for( int i = 0; i < n; i++ )
	{
		msurface_t *out = &mod->surfaces[i];
		out->info = &info[i]; info[i].surf = out;
		out->firstedge = df[i].firstedge;
		out->numedges = df[i].numedges;
		if( df[i].side ) out->flags |= SURF_PLANEBACK;
		out->plane = mod->planes + df[i].planenum;
		out->texinfo = mod->texinfo + df[i].texinfo;
		for( int j = 0; j < MAXLIGHTMAPS; j++ ) out->styles[j] = df[i].styles[j];

		const char *tn = out->texinfo->texture->name;
		if( !strncmp( tn, "sky", 3 )) out->flags |= SURF_DRAWSKY;
		if( xh_is_water_name( tn )) out->flags |= SURF_DRAWTURB;
		if( !strncmp( tn, "scroll", 6 )) out->flags |= SURF_CONVEYOR;
		if( out->texinfo->flags & TEX_SCROLL ) out->flags |= SURF_CONVEYOR;
		if( !strncmp( tn, "{scroll", 7 )) out->flags |= SURF_CONVEYOR | SURF_TRANSPARENT;
		if( tn[0] == '{' ) out->flags |= SURF_TRANSPARENT;
		if( out->texinfo->flags & TEX_SPECIAL ) out->flags |= SURF_DRAWTILED;

		xh_calc_bounds( mod, out );
		xh_calc_extents( mod, out );

		/* msurface_t::polys: the face's corners in edge order, as the GL renderers build them for every face
		 * (ref/gl/gl_rsurf.c GL_BuildPolygonFromSurface; positions only).  ref_soft builds none, and its decal placement
		 * then decides on an uninitialised count (ref/soft/r_decals.c:367, :578-583): with the polygon in place
		 * R_DecalVertsClip clips it against the decal as designed, and the oracle's answer is defined. */
		{
			glpoly2_t *poly = xh_alloc( sizeof( glpoly2_t ) + sizeof( float ) * VERTEXSIZE * out->numedges );
			poly->numverts = out->numedges;
			poly->flags = out->flags;
			for( int k = 0; k < out->numedges; k++ )
			{
				const int e = mod->surfedges[out->firstedge + k];
				const mvertex_t *v = e >= 0 ? &mod->vertexes[mod->edges16[e].v[0]] : &mod->vertexes[mod->edges16[-e].v[1]];
				memset( poly->verts[k], 0, sizeof( float ) * VERTEXSIZE );
				VectorCopy( v->position, poly->verts[k] );
			}
			out->polys = poly;
		}

		if( mod->lightdata && df[i].lightofs != -1 )
			out->samples = mod->lightdata + df[i].lightofs / 3;
		if( out->flags & SURF_DRAWTURB )
			xh.ref.GL_SubdivideSurface( mod, out );
	}

	/* marksurfaces (:2537) */
	const dmarkface_t *dmk = xh_lump( &b, LUMP_MARKSURFACES, sizeof( dmarkface_t ), &n );
	mod->nummarksurfaces = n;
	mod->marksurfaces = xh_alloc( sizeof( msurface_t * ) * ( n + 1 ));
	for( int i = 0; i < n; i++ )
	{
		if( dmk[i] >= mod->numsurfaces ) return -3;
		mod->marksurfaces[i] = mod->surfaces + dmk[i];
	}

	/* visibility (:3830) */
	{
		const dlump_t *l = &b.hdr->lumps[LUMP_VISIBILITY];
		if( l->filelen > 0 )
		{
			mod->visdata = xh_alloc( l->filelen );
			memcpy( mod->visdata, b.base + l->fileofs, l->filelen );
		}
	}

	/* leafs (:3546) */
	const dleaf_t *dl = xh_lump( &b, LUMP_LEAFS, sizeof( dleaf_t ), &n );
	int numleafs = n;
	int visclusters = mod->submodels[0].visleafs;
	xh.visbytes = ( visclusters + 7 ) >> 3;
	xh.refState.visbytes = xh.visbytes;
	xh.visscratch = xh_alloc( xh.visbytes + 16 );
	mod->leafs = xh_alloc( sizeof( mleaf_t ) * n );
	for( int i = 0; i < n; i++ )
	{
		mleaf_t *out = &mod->leafs[i];
		for( int j = 0; j < 3; j++ ) { out->minmaxs[j] = dl[i].mins[j]; out->minmaxs[j + 3] = dl[i].maxs[j]; }
		out->contents = dl[i].contents;
		out->firstmarksurface = mod->marksurfaces + dl[i].firstmarksurface;
		out->nummarksurfaces = dl[i].nummarksurfaces;
		memcpy( out->ambient_sound_level, dl[i].ambient_level, 4 );
		out->cluster = i - 1;
		if( out->cluster >= visclusters ) out->cluster = -1;
		int p = dl[i].visofs;
		out->compressed_vis = ( p == -1 || !mod->visdata ) ? NULL : mod->visdata + p;
		if( out->contents != CONTENTS_EMPTY )
			for( int j = 0; j < out->nummarksurfaces; j++ )
				out->firstmarksurface[j]->flags |= SURF_UNDERWATER;
	}
	if( mod->leafs[0].contents != CONTENTS_SOLID ) return -4;

	/* nodes (:3438) */
	const dnode_t *dn = xh_lump( &b, LUMP_NODES, sizeof( dnode_t ), &n );
	mod->numnodes = n;
	mod->nodes = xh_alloc( sizeof( mnode_t ) * n );
	for( int i = 0; i < n; i++ )
	{
		mnode_t *out = &mod->nodes[i];
		for( int j = 0; j < 3; j++ ) { out->minmaxs[j] = dn[i].mins[j]; out->minmaxs[j + 3] = dn[i].maxs[j]; }
		out->plane = mod->planes + dn[i].planenum;
		out->firstsurface_0 = dn[i].firstface;
		out->numsurfaces_0 = dn[i].numfaces;
		for( int j = 0; j < 2; j++ )
		{
			int p = dn[i].children[j];
			out->children_[j] = p >= 0 ? mod->nodes + p : (mnode_t *)( mod->leafs + ( -1 - p ));
		}
	}
	xh_set_parent( mod, mod->nodes, NULL );

	/* clipnodes: hull 0 is the render BSP itself (Mod_MakeHull0); brush entities only need
	 * hulls[0].firstclipnode == headnode (ref/soft/r_main.c:684 R_FindTopnode, r_bsp.c) */
	mod->hulls[0].firstclipnode = mod->submodels[0].headnode[0];

	/* submodels (:2110 Mod_SetupSubmodels): models[1] = world, models[1+i] = "*i" */
	xh.submodels = xh_alloc( sizeof( model_t ) * mod->numsubmodels );
	for( int i = 0; i < mod->numsubmodels; i++ )
	{
		dmodel_t *bm = &mod->submodels[i];
		model_t *sm = i == 0 ? mod : &xh.submodels[i];
		if( i != 0 )
		{
			*sm = *mod;
			snprintf( sm->name, sizeof( sm->name ), "*%i", i );
			sm->mempool = 0;
		}
		sm->hulls[0].firstclipnode = bm->headnode[0];
		sm->firstmodelsurface = bm->firstface;
		sm->nummodelsurfaces = bm->numfaces;
		for( int j = 0; j < 3; j++ ) { sm->mins[j] = bm->mins[j]; sm->maxs[j] = bm->maxs[j]; }
		{
			float r = 0;
			for( int j = 0; j < 3; j++ )
			{
				float a = fabsf( sm->mins[j] ), c = fabsf( sm->maxs[j] );
				float m = a > c ? a : c;
				r += m * m;
			}
			sm->radius = sqrtf( r );
		}
		sm->numleafs = bm->visleafs;
		sm->numframes = 2;
		int keep = mod->flags & MODEL_COLORED_LIGHTING;
		sm->flags = keep;
		if( i == 0 ) sm->flags |= MODEL_WORLD;
		for( int j = 0; i != 0 && j < sm->nummodelsurfaces; j++ )
		{
			msurface_t *s = mod->surfaces + sm->firstmodelsurface + j;
			if( s->flags & SURF_CONVEYOR ) sm->flags |= MODEL_CONVEYOR;
			if( s->flags & SURF_TRANSPARENT ) sm->flags |= MODEL_TRANSPARENT;
			if( s->flags & SURF_DRAWTURB ) sm->flags |= MODEL_LIQUID;
		}
		if( 1 + i <= MAX_MODELS ) xh.cl.models[1 + i] = sm;
	}
	mod->numleafs = mod->submodels[0].visleafs;
	(void)numleafs;
	xh.cl.nummodels = 1 + mod->numsubmodels;

	/* world entity: CL_GetEntityByIndex(0) must be a zeroed entity whose model is the world (r_bsp.c:633) */
	memset( &xh.entities[0], 0, sizeof( cl_entity_t ));
	xh.entities[0].model = mod;
	xh.entities[0].curstate.modelindex = 1;
	xh.num_entities = 1;

	for( int i = 0; i < mod->numsubmodels; i++ )
		xh.ref.Mod_ProcessRenderData( i == 0 ? mod : &xh.submodels[i], true, data, len );
	xh.ref.R_NewMap();
	xh_run_lightstyles();
	return 0;
}

void *xh_world_model( void ) { return xh.world; }

int xh_world_info( int *out, int n )
{
	model_t *m = xh.world;
	if( !m ) return 0;
	int v[] = { m->numsurfaces, m->numnodes, m->submodels[0].visleafs + 1, m->numedges, m->numvertexes,
		m->numsurfedges, m->nummarksurfaces, m->numplanes, m->numtextures, m->numsubmodels, (int)xh.visbytes };
	int c = (int)( sizeof( v ) / sizeof( v[0] ));
	for( int i = 0; i < c && i < n; i++ ) out[i] = v[i];
	return c;
}

int xh_surface_info( int surf, int *out, int n )
{
	model_t *m = xh.world;
	if( !m || surf < 0 || surf >= m->numsurfaces ) return 0;
	msurface_t *s = &m->surfaces[surf];
	int v[] = { s->flags, s->extents[0], s->extents[1], s->info->lightextents[0], s->info->lightextents[1],
		s->styles[0], s->styles[1], s->styles[2], s->styles[3], s->texinfo->texture->gl_texturenum,
		s->texturemins[0], s->texturemins[1], s->numedges, s->samples ? 1 : 0 };
	int c = (int)( sizeof( v ) / sizeof( v[0] ));
	for( int i = 0; i < c && i < n; i++ ) out[i] = v[i];
	return c;
}
