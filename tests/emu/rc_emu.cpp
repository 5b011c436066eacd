/*
 * rc_emu.cpp - DEBUG HARNESS, NOT A PRODUCT PATH.
 *
 * Implements the rc_dev.h entry points on the CPU by calling the very same scalar stage functions
 * (xash3d-fwgs_b200/csrc/rc_core.h) the CUDA kernels call, one item at a time.  Linked with rc_host.c into
 * tests/emu/libref_cuda_emu.so it lets the integer/float restatement of the ref_soft frame path be
 * checked against the oracle (oracle/_ref/libref_soft.so) on a machine without a GPU.  libref_cuda.so
 * never contains or loads this file; the GPU tests (-m gpu) exercise the kernels themselves.
 */
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "rc_dev.h"
#include "rc_core.h"
#include "rc_tri.h"

namespace {
struct Emu
{
	rc_scene_t sc{};
	std::vector<rc_plane_t> planes; std::vector<rc_node_t> nodes; std::vector<rc_leaf_t> leafs;
	std::vector<rc_surface_t> surfaces; std::vector<int32_t> marks, surfedges; std::vector<uint16_t> edges;
	std::vector<float> verts; std::vector<uint8_t> visdata, lightdata;
	std::vector<rc_texture_t> textures; std::vector<uint16_t> texels, colormap, lightgamma, cache;
	std::vector<int32_t> sintable, sedgepartner; std::vector<float> sedgeverts;
	std::vector<uint16_t> alphamap; std::vector<uint8_t> addmap, modmap;
	std::vector<rc_item_t> items; std::vector<rc_corner_t> corners; std::vector<rc_sortent_t> sortents;
	std::vector<rc_glow_t> glow; int nglow = 0;
	std::vector<rc_bent_t> bents;
	std::vector<uint8_t> sstatic; std::vector<rc_spart_t> parts; std::vector<float> bones, blight, mvpos; std::vector<int32_t> mlight;
	int32_t styles[256];
	struct Key { int32_t texnum; int32_t light[4]; bool valid; uint32_t decalgen; };
	std::vector<rc_ddecal_t> decals; std::vector<int32_t> decal_first; std::vector<uint32_t> decal_gen;
	std::vector<Key> cachekey;
	uint64_t counters[RC_CTR_COUNT];
	std::vector<std::vector<rc_surf_t>> vsurfs; std::vector<std::vector<rc_edge_t>> vedges;
	uint64_t launches = 0;
	char err[256] = "";
} E;

void fix_pointers()
{
	E.sc.planes = E.planes.data(); E.sc.nodes = E.nodes.data(); E.sc.leafs = E.leafs.data();
	E.sc.surfaces = E.surfaces.data(); E.sc.marksurfaces = E.marks.data(); E.sc.surfedges = E.surfedges.data();
	E.sc.edges = E.edges.data(); E.sc.verts = E.verts.data(); E.sc.visdata = E.visdata.data();
	E.sc.lightdata = E.lightdata.data(); E.sc.textures = E.textures.data(); E.sc.texels = E.texels.data();
	E.sc.colormap = E.colormap.data(); E.sc.lightgamma = E.lightgamma.data(); E.sc.sintable = E.sintable.data();
	E.sc.cache = E.cache.data();
	E.sc.numtextures = (int32_t)E.textures.size();
	E.sc.sedgeverts = E.sedgeverts.empty() ? nullptr : E.sedgeverts.data(); E.sc.sedgepartner = E.sedgepartner.empty() ? nullptr : E.sedgepartner.data();
}

void build_block( int s, int mip )
{
	const rc_surface_t *sf = &E.surfaces[s];
	const rc_texture_t *tex = &E.textures[sf->texnum];
	int smax = sf->lightextents[0] / sf->sample_size + 1, tmax = sf->lightextents[1] / sf->sample_size + 1;
	std::vector<uint32_t> bl( smax * tmax );
	for( int i = 0; i < smax * tmax; i++ )
		bl[i] = rc_blocklight( &E.sc, sf, E.styles, smax * tmax, i, 0u );
	int w = sf->lightextents[0] >> mip, h = sf->lightextents[1] >> mip;
	uint16_t *dst = E.cache.data() + sf->cache_ofs[mip];
	for( int y = 0; y < h; y++ )
		for( int x = 0; x < w; x++ )
			dst[y * w + x] = rc_cache_texel( &E.sc, sf, tex, mip, bl.data(), smax, x, y );
	if( !E.decal_first.empty() && E.decal_first[2 * s + 1] > 0 )
	{
		/* R_DrawSurfaceDecals over the finished block */
		rc_blendluts_t L = { E.colormap.data(), E.alphamap.data(), E.addmap.data(), E.modmap.data() };
		const rc_ddecal_t *dl = E.decals.data() + E.decal_first[2 * s];
		for( int y = 0; y < h; y++ )
			for( int x = 0; x < w; x++ )
			{
				uint32_t px = dst[y * w + x];
				if( rc_decal_texel( &L, E.texels.data(), dl, E.decal_first[2 * s + 1], mip, x, y, &px )) dst[y * w + x] = (uint16_t)px;
			}
	}
	Emu::Key &k = E.cachekey[s * 4 + mip];
	k.valid = true; k.texnum = sf->texnum; k.decalgen = E.decal_gen.empty() ? 0u : E.decal_gen[s];
	for( int j = 0; j < 4; j++ ) k.light[j] = sf->styles[j] != 255 ? E.styles[sf->styles[j]] : 0;
	E.counters[RC_CTR_CACHE_BLOCKS]++; E.counters[RC_CTR_CACHE_TEXELS] += (uint64_t)w * h;
}

bool block_stale( int s, int mip )
{
	const rc_surface_t *sf = &E.surfaces[s];
	const Emu::Key &k = E.cachekey[s * 4 + mip];
	if( !k.valid || k.texnum != sf->texnum || k.decalgen != ( E.decal_gen.empty() ? 0u : E.decal_gen[s] )) return true;
	for( int j = 0; j < 4 && sf->styles[j] != 255; j++ )
		if( k.light[j] != E.styles[sf->styles[j]] ) return true;
	return false;
}
}

extern "C" {

int rc_dev_init( int ) { for( int i = 0; i < 256; i++ ) E.styles[i] = 256; return 0; }
void rc_dev_shutdown( void ) {}
const char *rc_dev_last_error( void ) { return E.err; }
const char *rc_dev_name( void ) { return "cpu-emulation (debug harness)"; }
uint64_t rc_dev_launch_count( void ) { return 0; }
uint64_t rc_dev_h2d_bytes( int ) { return 0; }

int rc_dev_upload_luts( const uint16_t *colormap, const uint16_t *lightgamma, const int32_t *sintable )
{
	E.colormap.assign( colormap, colormap + 32 * 8192 );
	E.lightgamma.assign( lightgamma, lightgamma + 1024 );
	E.sintable.assign( sintable, sintable + 1280 );
	fix_pointers();
	return 0;
}

int rc_dev_upload_textures( const rc_texture_t *tex, const rc_texsrc_t *src, int ntex, const uint32_t *rgba, size_t, size_t ntexels )
{
	E.textures.assign( tex, tex + ntex );
	E.texels.assign( ntexels, 0 );
	/* k_tex_encode, sequentially: every plane of every texture, texel by texel */
	for( int i = 0; i < ntex; i++ )
	{
		const rc_texture_t &t = tex[i];
		if( t.ofs[0] < 0 ) continue;
		for( int plane = 0; plane < 5; plane++ )
		{
			const int level = plane == 4 ? 0 : plane;
			const int32_t ofs = plane == 4 ? t.alpha_ofs : t.ofs[plane];
			if( ofs < 0 ) continue;
			const int32_t n = std::max( 1, t.width >> level ) * std::max( 1, t.height >> level );
			for( int32_t k = 0; k < n; k++ )
				E.texels[(size_t)ofs + k] = rc_tex_encode( rgba + src[i].rgba_ofs, t.width, t.height, src[i].flags, plane, k );
		}
	}
	fix_pointers();
	return rc_dev_flush_cache();
}

int rc_dev_upload_scene( const rc_scene_t *sc, size_t cache_texels )
{
	E.sc = *sc;
	E.planes.assign( sc->planes, sc->planes + sc->numplanes );
	E.nodes.assign( sc->nodes, sc->nodes + sc->numnodes );
	E.leafs.assign( sc->leafs, sc->leafs + sc->numleafs );
	E.surfaces.assign( sc->surfaces, sc->surfaces + sc->numsurfaces );
	E.marks.assign( sc->marksurfaces, sc->marksurfaces + sc->nummarks );
	E.surfedges.assign( sc->surfedges, sc->surfedges + sc->numsurfedges );
	E.edges.assign( sc->edges, sc->edges + 2 * sc->numedges );
	E.verts.assign( sc->verts, sc->verts + 3 * sc->numverts );
	E.visdata.assign( sc->visdata, sc->visdata + sc->visdata_len );
	E.lightdata.assign( sc->lightdata, sc->lightdata + sc->lightdata_len );
	E.cache.assign( cache_texels, 0 );
	E.cachekey.assign( (size_t)sc->numsurfaces * 4, Emu::Key{} );
	E.decals.clear(); E.decal_first.clear(); E.decal_gen.clear();
	fix_pointers();
	E.sedgeverts.assign( (size_t)sc->numsurfedges * 6 + 6, 0.0f ); E.sedgepartner.assign( (size_t)sc->numsurfedges + 1, -1 );
	rc_build_sedge_tables( &E.sc, E.sedgeverts.data(), E.sedgepartner.data());
	fix_pointers();
	return 0;
}

int rc_dev_update_surface_texnums( const int32_t *texnum, int n )
{
	for( int i = 0; i < n && i < (int)E.surfaces.size(); i++ ) E.surfaces[i].texnum = texnum[i];
	return 0;
}

int rc_dev_upload_blendluts( const uint16_t *alphamap, const uint8_t *addmap, const uint8_t *modmap )
{
	E.alphamap.assign( alphamap, alphamap + 3 * 1024 * 256 );
	E.addmap.assign( addmap, addmap + 65536 ); E.modmap.assign( modmap, modmap + 65536 );
	return 0;
}

int rc_dev_set_items( const rc_item_t *items, int nitems, const rc_corner_t *corners, int ncorners, const rc_sortent_t *sortents, int nsortents )
{
	E.items.assign( items, items + nitems ); E.corners.assign( corners, corners + ncorners ); E.sortents.assign( sortents, sortents + nsortents );
	return 0;
}

int rc_dev_upload_studio_static( const void *arena, size_t bytes )
{
	E.sstatic.assign( (const uint8_t *)arena, (const uint8_t *)arena + bytes );
	return 0;
}

/* the batch's studio parts: vertices through their bones, one light level per normal (k_studio_points on the device) */
int rc_dev_set_studio( const rc_spart_t *parts, int nparts, const float *bones, int nbones, const float *blightvecs, int totalverts, int totalnorms )
{
	E.parts.assign( parts, parts + nparts );
	E.bones.assign( bones, bones + (size_t)nbones * 12 ); E.blight.assign( blightvecs, blightvecs + (size_t)nbones * 3 );
	E.mvpos.assign( (size_t)totalverts * 3 + 3, 0.0f ); E.mlight.assign( (size_t)totalnorms + 1, 0 );
	const float *sf = (const float *)E.sstatic.data();
	for( const rc_spart_t &pt : E.parts )
	{
		for( int i = 0; i < pt.numverts; i++ )
			rc_studio_vertex( &E.bones[(size_t)( pt.bones + E.sstatic[pt.svertbone + i] ) * 12], sf + pt.sverts + 3 * i, &E.mvpos[(size_t)( pt.vert_base + i ) * 3] );
		for( int k = 0; k < pt.numnorms; k++ )
			E.mlight[pt.light_base + k] = rc_studio_light( &pt, E.sstatic[pt.snormflags + k], sf + pt.snorms + 3 * k,
				&E.blight[(size_t)( pt.blightvec + E.sstatic[pt.snormbone + k] ) * 3], E.lightgamma.data());
	}
	return 0;
}

/* one polygon fan of the triangle rasteriser, scanline by scanline (what the kernels do per row / per pixel) */
/* RC_EMU_FASTKEYS=1: the leading opaque run is drawn the way k_tris_fast / k_tri_resolve do (rc_k_tris.cuh): a per-pixel
 * maximum of (z << 32 | submission number), then every pixel shaded from the triangle its key names */
struct emu_keyed_t { rc_trisetup_t ts; const rc_item_t *it; };
static std::vector<unsigned long long> K_keys;
static std::vector<emu_keyed_t> K_tris;
static bool K_active = false;

static void emu_draw_poly( const rc_item_t *it, const rc_fv_t *poly, int k, uint16_t *color, int16_t *zb, int width )
{
	rc_blendluts_t L = { E.colormap.data(), E.alphamap.data(), E.addmap.data(), E.modmap.data() };
	const uint16_t *skin = E.texels.data() + it->tex_ofs;
	for( int i = 1; i < k - 1; i++ )
	{
		rc_trisetup_t ts;
		if( !rc_tri_setup( &poly[0], &poly[i], &poly[i + 1], it->skinwidth, it->skinheight, &ts ))
			continue;
		E.counters[RC_CTR_TRIS]++;
		int32_t y0;
		const int32_t nrows = rc_tri_rows( &ts, &y0 );
		if( K_active )
		{
			K_tris.push_back( { ts, it } );
			const unsigned long long id = K_tris.size();
			for( int32_t r = 0; r < nrows; r++ )
			{
				rc_trirow_t row;
				rc_tri_row( &ts, r, &row );
				if( row.count <= 0 || !rc_tri_row_inbounds( &ts, &row ))
					continue;
				for( int32_t p = 0; p < row.count; p++ )
				{
					const size_t px = (size_t)( y0 + r ) * width + row.x + p;
					const int32_t zi = (int32_t)((uint32_t)row.zi + (uint32_t)p * (uint32_t)ts.zistepx );
					const unsigned long long key = ((unsigned long long)(uint32_t)(( zi >> 16 ) + 32768 ) << 32 ) | id;
					if( key > K_keys[px] ) K_keys[px] = key;
					E.counters[RC_CTR_TRI_PIXELS]++;
				}
			}
			continue;
		}
		for( int32_t r = 0; r < nrows; r++ )
		{
			rc_trirow_t row;
			rc_tri_row( &ts, r, &row );
			if( row.count <= 0 || !rc_tri_row_inbounds( &ts, &row ))
				continue;
			for( int32_t p = 0; p < row.count; p++ )
			{
				const size_t px = (size_t)( y0 + r ) * width + row.x + p;
				const uint32_t src = skin[rc_tri_texel( &ts, &row, p )];
				const int32_t light = (int32_t)((uint32_t)row.l + (uint32_t)p * (uint32_t)ts.lstepx );
				const int32_t zi = (int32_t)((uint32_t)row.zi + (uint32_t)p * (uint32_t)ts.zistepx );
				rc_tri_shade( &L, it->drawer, (uint32_t)it->color, (uint32_t)it->alpha, src, light, zi, &color[px], &zb[px] );
				E.counters[RC_CTR_TRI_PIXELS]++;
			}
		}
	}
}

static void emu_keys_resolve( uint16_t *color, int16_t *zb, int width, int height )
{
	for( int y = 0; y < height; y++ )
		for( int x = 0; x < width; x++ )
		{
			const size_t px = (size_t)y * width + x;
			const uint32_t id = (uint32_t)K_keys[px];
			if( !id ) continue;
			const emu_keyed_t &kt = K_tris[id - 1];
			int32_t y0;
			rc_tri_rows( &kt.ts, &y0 );
			rc_trirow_t row;
			rc_tri_row( &kt.ts, y - y0, &row );
			const int32_t p = x - row.x;
			const uint32_t src = E.texels[kt.it->tex_ofs + rc_tri_texel( &kt.ts, &row, p )];
			const uint32_t light = (uint32_t)row.l + (uint32_t)p * (uint32_t)kt.ts.lstepx;
			color[px] = (uint16_t)( E.colormap[( src >> 3 ) | (( light & 0x1F00u ) << 5 )] | ( src & 7 ));
			zb[px] = (int16_t)( (int32_t)( K_keys[px] >> 32 ) - 32768 );
		}
	K_active = false;
}

/* what a translucent brush entity needs of the view's world pass */
struct emu_viewctx_t { const uint32_t *nodevis, *leafvis; const int32_t *leafbase, *leafkey; const uint16_t *candof; };

/* R_DrawBrushModel: the entity's own edge frame, spans per scanline, D_AlphaSurf's drawers (k_tbrush_spans + k_tris_ordered) */
static void emu_draw_tbrush( const rc_view_t *vw, const emu_viewctx_t &cx, int32_t ent, uint16_t *color, int16_t *zb, int width )
{
	const rc_scene_t *sc = &E.sc;
	const rc_bent_t *be = &E.bents[ent];
	rc_blendluts_t L = { E.colormap.data(), E.alphamap.data(), E.addmap.data(), E.modmap.data() };
	std::vector<rc_surf_t> surfs( RC_TB_MAXSURFS );
	std::vector<rc_edge_t> edges( RC_TB_MAXEDGES );
	int32_t nedges = 0, ovf = 0;
	const int32_t ns = rc_tbrush_pieces( sc, vw, E.bents.data(), (int32_t)E.bents.size(), ent, cx.nodevis, cx.leafvis, cx.leafbase, cx.leafkey, cx.candof,
		surfs.data(), edges.data(), &nedges, nullptr, &ovf );
	if( ovf ) E.counters[RC_CTR_OVERFLOW_BMODEL]++;
	if( !ns ) return;
	E.counters[RC_CTR_BMODEL_SURFS] += ns;
	for( int32_t k = 0; k < ns; k++ )
	{
		const rc_surface_t *sf = &E.surfaces[surfs[k].surf];
		if( !( sf->flags & ( RC_SURF_DRAWSKY | RC_SURF_DRAWTURB )) && sf->cache_ofs[0] >= 0 && block_stale( surfs[k].surf, 0 ))
			build_block( surfs[k].surf, 0 );
	}
	const int32_t *turb = E.sintable.data() + vw->turb_phase;
	const int x0 = vw->vrect[0], y0 = vw->vrect[1], vw_w = vw->vrect[2], vw_h = vw->vrect[3];
	const int32_t head_iu = x0, tail_u = (( x0 + vw_w ) << 20 ) + 0xFFFFF;
	for( int y = y0; y < y0 + vw_h; y++ )
	{
		rc_span_t spans[RC_TB_LINESPANS];
		const int32_t nsp = rc_tbrush_line( surfs.data(), edges.data(), nedges, y, head_iu, tail_u, spans, RC_TB_LINESPANS, &ovf );
		for( int32_t si = 0; si < nsp; si++ )
		{
			const rc_span_t &sp = spans[si];
			if( !sp.surf ) continue;
			const rc_surf_t *sf = &surfs[sp.surf - 1];
			const int kind = rc_span_kind( sf );
			if( kind == RC_KIND_SKY ) continue;
			if( kind == RC_KIND_TURB && be->alpha == 0 ) continue;
			const int nch = rc_span_nchunks( kind, sp.count ), cpx = kind == RC_KIND_TURB ? 16 : 8;
			for( int j = 0; j < nch; j++ )
			{
				rc_chunk_t c;
				rc_chunk_setup( vw, sf, vw->sky_color, sp.u, y, sp.count, j, &c );
				const int n = std::min( cpx, sp.count - j * cpx );
				for( int i = 0; i < n; i++ )
				{
					const size_t p = (size_t)y * width + c.u0 + i;
					const int32_t ti = rc_pixel_texel( &c, i, turb );
					const uint32_t texel = kind == RC_KIND_TURB ? E.texels[ti] : E.cache[ti];
					const int32_t izi = (int32_t)((uint32_t)c.izi + (uint32_t)i * (uint32_t)c.izistep );
					rc_alpha_pixel( &L, be->trans, (uint32_t)std::min( be->alpha, 7 ), kind == RC_KIND_TURB, texel, izi, &color[p], &zb[p] );
				}
			}
		}
	}
	if( ovf ) E.counters[RC_CTR_OVERFLOW_EVENTS]++;
}

static void emu_draw_items( const rc_view_t *vw, int vi, const emu_viewctx_t &cx, uint16_t *color, int16_t *zb, int width, int height )
{
	if( E.items.empty() || !color || !zb )
		return;
	/* item order: fixed items as listed; the translucent entities in R_TransEntityCompare's order for this view */
	std::vector<int> order;
	std::vector<int> ents( E.sortents.size());
	for( size_t i = 0; i < ents.size(); i++ ) ents[i] = (int)i;
	std::vector<float> dist( E.sortents.size());
	for( size_t i = 0; i < ents.size(); i++ )
	{
		const rc_sortent_t &se = E.sortents[i];
		float v[3] = { vw->origin[0] - se.center[0], vw->origin[1] - se.center[1], vw->origin[2] - se.center[2] };
		dist[i] = se.use_dist ? rc_dot3( v, v ) : 1000000000.0f;
	}
	std::stable_sort( ents.begin(), ents.end(), [&]( int a, int b ) {
		if( dist[a] != dist[b] ) return dist[a] > dist[b];
		return E.sortents[a].rank < E.sortents[b].rank; } );
	bool trans_done = false;
	for( size_t i = 0; i < E.items.size(); i++ )
	{
		if( E.items[i].sortent < 0 ) { order.push_back( (int)i ); continue; }
		if( trans_done ) continue;
		trans_done = true;
		for( int e : ents )
			for( int j = 0; j < E.sortents[e].nitems; j++ ) order.push_back( E.sortents[e].first_item + j );
	}
	int nfast = 0;
	if( getenv( "RC_EMU_FASTKEYS" ))
	{
		int tris = 0;
		for( size_t i = 0; i < E.items.size() && E.items[i].drawer == RC_DRAW_FILL && E.items[i].sortent < 0; i++ ) { nfast++; tris += E.items[i].count; }
		if( tris < 64 ) nfast = 0;
	}
	if( nfast )
	{
		K_keys.resize( (size_t)width * height ); K_tris.clear(); K_active = true;
		for( size_t i = 0; i < K_keys.size(); i++ ) K_keys[i] = (unsigned long long)(uint32_t)( (int32_t)zb[i] + 32768 ) << 32;
	}
	for( int idx : order )
	{
		if( K_active && idx >= nfast ) emu_keys_resolve( color, zb, width, height );
		const rc_item_t *it = &E.items[idx];
		if( it->tex_ofs < 0 ) continue;
		if( it->kind == RC_ITEM_BRUSH ) { emu_draw_tbrush( vw, cx, it->first, color, zb, width ); continue; }
		rc_fv_t poly[8];
		if( rc_item_is_quad( it ))
		{
			rc_corner_t q[4];
			rc_item_t gi;
			if( it->glow >= 0 )
			{
				/* a glow sprite: this view's verdict, size and colour */
				const rc_glow_t &g = E.glow[(size_t)vi * E.nglow + it->glow];
				if( !g.visible ) continue;
				gi = *it; gi.scale = g.scale; gi.light = g.light; gi.color = g.color; gi.alpha = g.alpha;
				it = &gi;
			}
			if( !rc_quad_corners( vw, it, q )) continue;
			int k = rc_view_triangle( vw, &q[0], &q[1], &q[2], poly );
			if( k ) emu_draw_poly( it, poly, k, color, zb, width );
			k = rc_view_triangle( vw, &q[0], &q[2], &q[3], poly );
			if( k ) emu_draw_poly( it, poly, k, color, zb, width );
		}
		else if( it->kind == RC_ITEM_MESH )
		{
			const rc_mcorner_t *mc = (const rc_mcorner_t *)E.sstatic.data() + it->first;
			for( int t = 0; t < it->count; t++ )
			{
				rc_corner_t c[3];
				for( int q = 0; q < 3; q++ )
				{
					const rc_mcorner_t &m = mc[3 * t + q];
					const float *p = &E.mvpos[(size_t)( it->vert_base + m.vert ) * 3];
					c[q].xyz[0] = p[0]; c[q].xyz[1] = p[1]; c[q].xyz[2] = p[2];
					c[q].l = E.mlight[it->light_base + m.norm]; c[q].s = m.s; c[q].t = m.t;
					if( it->chrome )
					{
						const rc_spart_t *pt = &E.parts[it->part];
						int32_t cs, ct;
						rc_studio_chrome_st( vw, E.bones.data() + (size_t)( pt->bones + E.sstatic[pt->snormbone + m.norm] ) * 12,
							(const float *)E.sstatic.data() + pt->snorms + 3 * m.norm, it->skinwidth, it->skinheight, &cs, &ct );
						c[q].s = (int16_t)cs; c[q].t = (int16_t)ct;
					}
				}
				int k = rc_view_triangle( vw, &c[0], &c[1], &c[2], poly );
				if( k ) emu_draw_poly( it, poly, k, color, zb, width );
			}
		}
		else if( it->kind == RC_ITEM_TRIS )
		{
			for( int t = 0; t < it->count; t++ )
			{
				const rc_corner_t *c = &E.corners[it->first + 3 * t];
				int k = rc_view_triangle( vw, &c[0], &c[1], &c[2], poly );
				if( k ) emu_draw_poly( it, poly, k, color, zb, width );
			}
		}
	}
	if( K_active ) emu_keys_resolve( color, zb, width, height );
}

int rc_dev_wait( void ) { return 0; }
int rc_dev_set_lightstyles( const int32_t *values ) { memcpy( E.styles, values, sizeof( E.styles )); return 0; }
int rc_dev_flush_cache( void ) { for( auto &k : E.cachekey ) k.valid = false; return 0; }

int rc_dev_render( const rc_view_t *views, int nviews, int width, int height, int clear, uint16_t *color_out, int16_t *z_out )
{
	const rc_scene_t *sc = &E.sc;
	memset( E.counters, 0, sizeof( E.counters ));
	E.vsurfs.assign( nviews, {} ); E.vedges.assign( nviews, {} );
	size_t px = (size_t)width * height;
	int stack_ints = 2 * ( sc->maxdepth + 4 ) + 8;

	for( int vi = 0; vi < nviews; vi++ )
	{
		const rc_view_t *vw = &views[vi];
		uint16_t *color = color_out ? color_out + px * vi : nullptr;
		int16_t *zb = z_out ? z_out + px * vi : nullptr;
		if( clear )
		{
			if( color ) memset( color, 0, px * 2 );
			if( zb ) memset( zb, 0, px * 2 );
		}
		/* A1 */
		std::vector<uint8_t> vis( sc->visbytes + 8 );
		std::vector<int32_t> stack( stack_ints + 64 );
		if( vw->pvs_radius < 0 ) std::fill( vis.begin(), vis.end(), 0xFF );
		else rc_fat_pvs( sc, vw->origin, vw->pvs_radius, vis.data(), stack.data(), stack_ints );
		std::vector<uint32_t> nodevis(( sc->numnodes + 31 ) / 32 + 1 ), leafmark(( sc->numleafs + 31 ) / 32 + 1 ), surfmark(( sc->numsurfaces + 31 ) / 32 + 1 );
		rc_mark_leaves( sc, vis.data(), nodevis.data(), leafmark.data());
		/* A2 */
		int32_t maxc = sc->numsurfaces;
		for( const rc_bent_t &b : E.bents ) maxc += b.numfaces * 64;
		std::vector<rc_cand_t> cands( maxc + 1 );
		std::vector<int32_t> leafbase( sc->numleafs + 1, -1 ), leafkey( sc->numleafs + 1, -1 );
		int32_t overflow = 0;
		int ncands = rc_walk_world( sc, vw, nodevis.data(), leafmark.data(), surfmark.data(), stack.data(), stack_ints, cands.data(), maxc, &overflow,
			E.bents.data(), (int32_t)E.bents.size(), leafbase.data(), leafkey.data());
		if( overflow ) E.counters[RC_CTR_OVERFLOW_STACK]++;
		E.counters[RC_CTR_CANDS] += ncands;
		/* A3 + A5 */
		std::vector<rc_surf_t> surfs( ncands + 1 );
		std::vector<uint16_t> candof( sc->numsurfaces + 2, 0 );
		for( int ci = 0; ci < ncands; ci++ ) if( cands[ci].surf >= 0 ) candof[cands[ci].surf] = (uint16_t)( ci + 1 );
		std::vector<rc_edge_t> edges;
		std::vector<rc_facerec_t> recs( ncands + 1 );
		for( int ci = 0; ci < ncands; ci++ )
		{
			if( cands[ci].surf < 0 ) { memset( &surfs[ci], 0, sizeof( rc_surf_t )); memset( &recs[ci], 0, sizeof( rc_facerec_t )); continue; } /* a brush entity's slot */
			/* as the kernels do it: k_faces begins every face and leaves its record; a face that has to inherit a clip
			 * point is finished afterwards (k_faces_fix) - the record of such a face keeps what its edge loop wrote */
			rc_edge_t local[256];
			rc_facestate_t fs; memset( &fs, 0, sizeof( fs ));
			fs.out = local; fs.maxout = 256; fs.surfnum = ci + 1;
			int ml, mr;
			const int need = rc_face_begin( sc, vw, cands.data(), candof.data(), ci, &fs, &ml, &mr );
			if( need )
			{
				rc_face_store( &fs, ml, mr, 1, &recs[ci] );
				rc_face_inherit( recs.data(), ci, need, &fs );
				rc_face_finish( sc, vw, &cands[ci], &fs, ml, mr, &surfs[ci] );
			}
			else
			{
				rc_face_finish( sc, vw, &cands[ci], &fs, ml, mr, &surfs[ci] );
				rc_face_store( &fs, ml, mr, 0, &recs[ci] );
			}
			for( int k = 0; k < fs.nout && k < 256; k++ ) edges.push_back( local[k] );
			if( surfs[ci].emitted )
			{
				E.counters[RC_CTR_SURFS]++;
				const rc_surface_t *sf = &E.surfaces[cands[ci].surf];
				/* A6: the block this view samples must be current */
				if( !( sf->flags & ( RC_SURF_DRAWSKY | RC_SURF_DRAWTURB )) && sf->cache_ofs[surfs[ci].miplevel] >= 0
					&& block_stale( cands[ci].surf, surfs[ci].miplevel ))
					build_block( cands[ci].surf, surfs[ci].miplevel );
			}
		}
		/* A13: the opaque brush entities, face by face (k_bmodels) */
		if( !E.bents.empty())
		{
			std::vector<rc_edge_t> bedges( 64 * 1024 );
			int32_t nbe = 0;
			rc_bmout_t out;
			out.surfs = surfs.data(); out.max_cands = ncands;
			out.edges = bedges.data(); out.max_edges = (int32_t)bedges.size(); out.nedges = &nbe;
			out.leafbase = leafbase.data(); out.candof = candof.data(); out.cache_need = nullptr; out.leafkey = nullptr;
			out.nsurfs = 0; out.overflow = 0; out.noutmax = 0;
			for( size_t e = 0; e < E.bents.size(); e++ )
			{
				const rc_bent_t *be = &E.bents[e];
				if( be->trans ) continue; /* drawn in its turn among the sorted entities (emu_draw_tbrush) */
				const int32_t clipflags = rc_bmodel_check_bbox( vw, be->minmaxs );
				if( clipflags < 0 ) continue;
				const int32_t topnode = rc_find_topnode( sc, nodevis.data(), leafmark.data(), be->minmaxs );
				if( topnode == RC_TOPNODE_NONE ) continue;
				rc_view_t bv;
				rc_bmodel_view( vw, be, &bv );
				for( int f = 0; f < be->numfaces; f++ )
					rc_bmodel_face( sc, &bv, E.bents.data(), (int32_t)E.bents.size(), (int32_t)e, be->firstface + f, clipflags, topnode,
						nodevis.data(), leafmark.data(), &out );
			}
			for( int32_t k = 0; k < nbe && k < (int32_t)bedges.size(); k++ ) edges.push_back( bedges[k] );
			E.counters[RC_CTR_BMODEL_SURFS] += out.nsurfs; E.counters[RC_CTR_SURFS] += out.nsurfs;
			if( out.overflow ) E.counters[RC_CTR_OVERFLOW_BMODEL]++;
			for( int ci = 0; ci < ncands; ci++ )
				if( cands[ci].surf < 0 && surfs[ci].emitted )
				{
					const int32_t face = surfs[ci].surf;
					const rc_surface_t *sf = &E.surfaces[face];
					if( !( sf->flags & ( RC_SURF_DRAWSKY | RC_SURF_DRAWTURB )) && sf->cache_ofs[surfs[ci].miplevel] >= 0 && block_stale( face, surfs[ci].miplevel ))
						build_block( face, surfs[ci].miplevel );
				}
		}
		E.counters[RC_CTR_EDGES] += edges.size();
		/* A4 + A9..A11 per scanline */
		const int32_t *turb = E.sintable.data() + vw->turb_phase;
		int x0 = vw->vrect[0], y0 = vw->vrect[1], vw_w = vw->vrect[2], vw_h = vw->vrect[3];
		int32_t head_iu = x0, tail_u = (( x0 + vw_w ) << 20 ) + 0xFFFFF;
		std::vector<rc_event_t> ev; std::vector<rc_span_t> spans( edges.size() + 2 );
		std::vector<int32_t> stk( ncands + 2 ), cnt( ncands + 2 ), neg( 2 * 64 );
		for( int y = y0; y < y0 + vw_h; y++ )
		{
			ev.clear();
			for( const rc_edge_t &e : edges )
				if( rc_edge_active( &e, y ))
					ev.push_back( rc_event_t{ rc_edge_u_at( &e, y ), e.surfs } );
			std::sort( ev.begin(), ev.end(), []( const rc_event_t &a, const rc_event_t &b ) { return rc_event_key( a.u, a.surfs ) < rc_event_key( b.u, b.surfs ); } );
			int32_t ovf = 0;
			std::vector<unsigned long long> keys( ev.size());
			for( size_t k = 0; k < ev.size(); k++ ) keys[k] = rc_event_key( ev[k].u, ev[k].surfs );
			int ns = rc_generate_spans( keys.data(), (int)ev.size(), head_iu, tail_u, nullptr, stk.data(), cnt.data(), ncands + 1, neg.data(), 64,
				spans.data(), (int)spans.size(), &ovf, E.bents.empty() ? nullptr : surfs.data(), y );
			{
				/* the bit-set formulation the scanline kernel uses must give the same spans */
				std::vector<uint32_t> bits(( ncands + 32 ) / 32 + 1, 0u );
				std::vector<rc_span_t> spans2( spans.size());
				int32_t exc[32], ovf2 = 0;
				int ns2 = rc_generate_spans_bits( keys.data(), (int)ev.size(), head_iu, tail_u, bits.data(), (int)bits.size(), exc, 16,
					spans2.data(), (int)spans2.size(), &ovf2 );
				bool same = ns2 == ns && ovf2 == ovf;
				for( int k = 0; same && k < ns; k++ )
					same = spans[k].u == spans2[k].u && spans[k].count == spans2[k].count && spans[k].surf == spans2[k].surf;
				if( !same && !E.bents.empty())
					E.counters[RC_CTR_BMODEL_TIES]++; /* pieces of brush entities that share a leaf key overlapped: the 1/z order decided */
				else if( !same )
				{
					fprintf( stderr, "rc_emu: span replay mismatch view %d line %d (%d vs %d spans)\n", vi, y, ns, ns2 );
					E.counters[RC_CTR_OVERFLOW_CANDS]++;
				}
			}
			if( ovf ) E.counters[RC_CTR_OVERFLOW_EVENTS]++;
			E.counters[RC_CTR_SPANS] += ns;
			if( const char *dbg = getenv( "RC_EMU_DEBUG" ))
			{
				int dv, dx, dy;
				if( sscanf( dbg, "%d,%d,%d", &dv, &dx, &dy ) == 3 && dv == vi && dy == y )
				{
					for( size_t k = 0; k < ev.size(); k++ )
						fprintf( stderr, "ev %zu u=%d iu=%d frac=%05x trail=%u lead=%u\n", k, ev[k].u, ev[k].u >> 20, ev[k].u & 0xFFFFF, ev[k].surfs & 0xFFFF, ev[k].surfs >> 16 );
					for( int k = 0; k < ns; k++ )
						fprintf( stderr, "span u=%d count=%d surf=%d face=%d mip=%d\n", spans[k].u, spans[k].count, spans[k].surf, spans[k].surf ? surfs[spans[k].surf - 1].surf : -1, spans[k].surf ? surfs[spans[k].surf - 1].miplevel : -1 );
				}
			}
			for( int si = 0; si < ns; si++ )
			{
				const rc_span_t &sp = spans[si];
				if( !sp.surf ) continue; /* background: pixels keep their previous content */
				const rc_surf_t *sf = &surfs[sp.surf - 1];
				int kind = rc_span_kind( sf );
				int nch = rc_span_nchunks( kind, sp.count ), cpx = kind == RC_KIND_TURB ? 16 : 8;
				for( int j = 0; j < nch; j++ )
				{
					rc_chunk_t c;
					rc_chunk_setup( vw, sf, vw->sky_color, sp.u, y, sp.count, j, &c );
					int n = std::min( cpx, sp.count - j * cpx );
					for( int i = 0; i < n; i++ )
					{
						size_t p = (size_t)y * width + c.u0 + i;
						if( color )
						{
							if( kind == RC_KIND_SKY ) color[p] = (uint16_t)c.base;
							else
							{
								int32_t ti = ( kind == RC_KIND_SOLID && vw->texfilt ) ? rc_pixel_texel_filt( &c, i, n, sp.u, y ) : rc_pixel_texel( &c, i, turb );
								color[p] = kind == RC_KIND_TURB ? E.texels[ti] : E.cache[ti];
							}
						}
						if( zb ) zb[p] = rc_pixel_z( &c, i );
					}
				}
			}
		}
		{
			emu_viewctx_t cx = { nodevis.data(), leafmark.data(), leafbase.data(), leafkey.data(), candof.data() };
			emu_draw_items( vw, vi, cx, color, zb, width, height );
		}
		E.vsurfs[vi] = surfs; E.vedges[vi] = edges;
		E.counters[RC_CTR_VIEWS]++; E.counters[RC_CTR_PIXELS] += (uint64_t)vw_w * vw_h;
	}
	return 0;
}

int rc_dev_render_async( const rc_view_t *views, int nviews, int width, int height, int clear, uint16_t *color_out, int16_t *z_out )
{
	return rc_dev_render( views, nviews, width, height, clear, color_out, z_out ); /* nothing is asynchronous here */
}

/* dynamic lights are not part of the debug harness */
int rc_dev_set_dlights( const rc_dlight_t *, int n ) { return n ? -1 : 0; }

/* the display tail is not part of the debug harness (it keeps no frames) */
int rc_dev_set_screenmap( const uint32_t *, int ) { return 0; }
int rc_dev_set_decals( const rc_ddecal_t *recs, int nrecs, const int32_t *face_first_count, const uint32_t *face_gen, int nfaces )
{
	E.decals.assign( recs, recs + ( nrecs > 0 ? nrecs : 0 ));
	E.decal_first.assign( face_first_count, face_first_count + 2 * (size_t)nfaces );
	E.decal_gen.assign( face_gen, face_gen + nfaces );
	return 0;
}
int rc_dev_set_glow( const rc_glow_t *recs, int nglow, int nviews )
{
	E.nglow = nglow > 0 ? nglow : 0;
	E.glow.assign( recs, recs + (size_t)E.nglow * ( nviews > 0 ? nviews : 0 ));
	return 0;
}
int rc_dev_set_bents( const rc_bent_t *bents, int n ) { E.bents.assign( bents, bents + ( n > 0 ? n : 0 )); return 0; }
int rc_dev_draw2d( const rc_cmd2d_t *, int, int ) { snprintf( E.err, sizeof( E.err ), "HUD pass: not in the debug harness" ); return -1; }
int rc_dev_blit_static( const uint32_t *, int, const uint16_t *, size_t, void * ) { snprintf( E.err, sizeof( E.err ), "display tail: not in the debug harness" ); return -1; }
int rc_dev_blit( int, int, void *, size_t, float * ) { snprintf( E.err, sizeof( E.err ), "display tail: not in the debug harness" ); return -1; }

int rc_dev_bench_resident( int, int, float *ms_total, float *ms_stage )
{
	if( ms_total ) *ms_total = 0;
	if( ms_stage ) memset( ms_stage, 0, RC_BENCH_STAGES * sizeof( float ));
	return -1;
}

int rc_dev_read_cache_block( int surf, int mip, uint16_t *out, int maxtexels, int *width, int *height, int *valid )
{
	if( surf < 0 || surf >= (int)E.surfaces.size() || mip < 0 || mip > 3 ) return -1;
	const rc_surface_t *sf = &E.surfaces[surf];
	int w = sf->lightextents[0] >> mip, h = sf->lightextents[1] >> mip;
	*width = w; *height = h;
	*valid = sf->cache_ofs[mip] >= 0 && E.cachekey[surf * 4 + mip].valid && !block_stale( surf, mip );
	if( *valid && w * h <= maxtexels )
		memcpy( out, E.cache.data() + sf->cache_ofs[mip], (size_t)w * h * 2 );
	return 0;
}

int rc_dev_read_counters( uint64_t *out, int n ) { for( int i = 0; i < n && i < RC_CTR_COUNT; i++ ) out[i] = E.counters[i]; return 0; }

int rc_dev_read_view_surfs( int view, rc_surf_t *out, int max, int *n )
{
	if( view < 0 || view >= (int)E.vsurfs.size()) return -1;
	*n = (int)E.vsurfs[view].size() - 1;
	for( int i = 0; i < *n && i < max; i++ ) out[i] = E.vsurfs[view][i];
	return 0;
}

int rc_dev_read_view_edges( int view, rc_edge_t *out, int max, int *n )
{
	if( view < 0 || view >= (int)E.vedges.size()) return -1;
	*n = (int)E.vedges[view].size();
	for( int i = 0; i < *n && i < max; i++ ) out[i] = E.vedges[view][i];
	return 0;
}

}
