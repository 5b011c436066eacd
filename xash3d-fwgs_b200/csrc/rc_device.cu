/*
 * rc_device.cu - the sm_100a kernels of ref_cuda and the thin C layer (rc_dev.h) the plugin's C host
 * code calls.  One batch = N independent camera views of one world at one timestamp.
 *
 * Stage map (reference function -> kernel), all per batch:
 *   S0 k_viswalk      Mod_PointInLeaf/Mod_FatPVS/R_MarkLeaves + R_RecursiveWorldNode   one block per view (rc_k_vis.cuh)
 *   S1 k_faces        R_RenderFace/R_ClipEdge/R_EmitEdge + D_CalcGradients             one thread per (view, face)
 *   S2 k_cache_list   D_CacheSurface staleness test -> work list                        one thread per (face, mip)
 *      k_cache_light  R_BuildLightMap (+ R_AddDynamicLights), light-block source offsets   warp per listed (face, mip)
 *      k_cache_texels R_DrawSurface / R_DrawSurfaceBlock8_mipN                            CTA per 256 runs of 8 / 4 / 2 texels, flat over all items
 *      k_cache_build  the same for blocks that are not made of whole light blocks          persistent CTAs, one block per (face, mip)
 *      k_dlight_mark  R_PushDlights/R_MarkLights                                        one thread per face, once per light set
 *   S3 k_events/k_lines R_ScanEdges/R_GenerateSpans (rc_k_spans.cuh)                    lane per scanline / warp per scanline
 *      k_draw         D_DrawSpans16/D_DrawZSpans/Turbulent8/D_BackgroundSurf (rc_k_draw.cuh) one warp per group of scanlines
 *   S4 k_blit         R_BlitScreen with the R_BuildScreenMap table                      one thread per 8 pixels
 *
 * The per-item arithmetic is rc_core.h (shared with the CPU debug harness); this file is the parallel
 * decomposition, the memory layout and the stream plumbing.  Compiled with -fmad=false and IEEE
 * division so float set-up matches the oracle's x86-64 SSE build bit for bit.
 */
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include <map>
#include <vector>

#include "rc_dev.h"
#include "rc_core.h"
#include "rc_tri.h"

/* ---------------------------------------------------------------------------------------------- */
/* device-side records private to this file                                                          */

struct rc_cachekey_t
{
	int32_t texnum;
	int32_t light[RC_MAXLIGHTMAPS];
	int32_t valid;
	int32_t dlight; /* surfcache_t.dlight: built with dynamic lights, never reused (r_surf.c:1099,1149) */
	int32_t decalgen; /* the face's decal generation the block was built at (a new decal sets surf->dlightframe: every mip is rebuilt) */
}; /* 32 B: what D_CacheSurface compares (r_surf.c:1090-1120): texture, light style values */

struct rc_lineinfo_t { int32_t ofs, nsp; }; /* span slice of a scanline: first slot inside the view's arena, count */
#define RC_LINE_DONE 0x40000000 /* set in nsp by k_lines: the slice holds spans (before: nsp = number of events) */

struct rc_work_t
{
	const rc_view_t *views;
	int32_t nviews, width, height, clear;
	int32_t max_cands, max_edges; /* per-view capacities */
	int32_t visbytes_padded, nodewords, leafwords, surfwords, stack_ints;
	int32_t vis_stage, vis_smem; /* k_viswalk: VW_STAGE_* flags and dynamic shared memory bytes */
	int32_t   *walkstack; /* [nviews][stack_ints] k_viswalk scratch (rc_vis_scratch_ints) */
	rc_cand_t *cands;
	uint16_t  *candof;    /* [nviews][numsurfaces] candidate index + 1 of a face in its view, 0 = not listed */
	rc_facerec_t *facerecs; /* [nviews][max_cands] what every face left in r_rast.c's clip-point statics */
	uint8_t   *facewrote; /* [nviews][max_cands] rc_facerec_t.wrote, packed for the backward search */
	int32_t   *fix_list, *fix_count; /* faces that wait for an earlier face's clip point: {view, candidate, first edge slot} */
	int32_t    fix_cap;
	rc_surf_t *surfs;
	rc_edge_t *edges;
	int32_t   *ncands, *nedges;
	uint32_t  *cache_need;/* [numsurfaces * 4 bits] */
	rc_cachekey_t *cache_keys; /* [numsurfaces * 4] */
	int32_t   *cache_list;/* [numsurfaces * 4] work list of surf * 4 + mip */
	/* the flat path of the surface cache (k_cache_light / k_cache_texels): (face, mip) blocks made of whole light blocks */
	int32_t   *cache_flist;   /* [numsurfaces * 4] their slots */
	uint32_t  *cache_units;   /* [numsurfaces * 4 * 32] pieces of 256 runs: position in cache_flist << 6 | piece */
	uint32_t  *cache_lux;     /* [numsurfaces * 4][CACHE_FLUX] blocklights of the item (rc_blocklight) */
	int32_t   *cache_src;     /* [numsurfaces * 4][CACHE_FLUX] source offset of every light block (the reference's wrap rules) */
	int32_t   *cache_fn;      /* [0] items in cache_flist, [1] pieces in cache_units */
	int32_t   *cache_nlist, *cache_cursor;
	int32_t   *line_cursor; /* k_lines work cursors: heavy list, all lines */
	int32_t   *heavy_list, *heavy_count; /* [nviews * height] lines with more than LN_HEAVY events (k_events), their number */
	int32_t   *wide_list, *wide_count;   /* [nviews * height] lines for k_lines_wide (RC_LINE_WIDE), their number */
	const int32_t *styles;/* [256] g_lightstylevalue */
	const rc_dlight_t *dlights; /* active dynamic lights of the batch */
	const uint32_t *dlightbits; /* [numsurfaces] msurface_t.dlightbits (k_dlight_mark) */
	int32_t ndlights;
	uint32_t  *scanbits;  /* surface bit sets of k_spans when a view lists > 2015 faces, else NULL */
	int32_t    scanbits_words, nbands;
	uint16_t  *bandsurf;  /* [nviews][nbands][max_cands + 2] band-local surface number -> candidate index + 1 */
	int32_t   *bandnloc;  /* [nviews][nbands] surfaces present in the band */
	int32_t    blk_events;           /* capacity: event/span slots per band */
	int32_t    ev_ecap;              /* band edges k_events stages in shared memory */
	size_t     view_events;          /* nbands * blk_events */
	unsigned long long *events;      /* [nviews][nbands][blk_events] sort keys, then rc_span_t in place */
	rc_lineinfo_t *lineinfo;         /* [nviews][height] span slice of every scanline */
	uint16_t  *color;
	int16_t   *zbuf;
	unsigned long long *counters;
	/* entities (rc_k_tris.cuh) */
	const rc_item_t *items; const rc_corner_t *corners; const rc_sortent_t *sortents;
	int32_t    nitems, nsortents;
	const rc_glow_t *glow; int32_t nglow; /* kRenderGlow sprites: [view][sprite] records of this slice's views */
	/* decals (R_DrawSurfaceDecals): records grouped by face, per face {first, count}, per face the generation; null = none */
	const rc_ddecal_t *decals; const int32_t *decal_first; const uint32_t *decal_gen;
	int32_t   *item_order;  /* [nviews][nitems] */
	struct rc_poly_t *polys; /* [nviews][ntris] */
	const uint16_t *alphamap; const uint8_t *addmap, *modmap;
	int32_t    nfast_items, fast_tris;  /* the leading run of opaque items: k_tris_fast instead of k_tris_ordered */
	unsigned long long *trikeys;        /* [nviews][height][width] z << 32 | triangle number */
	int32_t   *pairs, *pair_count;      /* k_item_cull: the (view << 16 | item) pairs of the opaque run that survive the cull */
	int32_t   *big_list, *big_count; int32_t big_cap; /* triangles k_tris_fast hands to k_tris_fast_big: {view, key number} */
	/* studio models (k_studio_points) */
	const uint8_t *sstatic; const rc_spart_t *parts; const float *bones, *blight; float *mvpos; int32_t *mlight;
	/* brush entities (k_bmodels) */
	const rc_bent_t *bents; int32_t nbents; const int32_t *bfaces; int32_t nbfaces; /* bfaces: entity << 16 | face within it */
	int32_t   *leafbase;                /* [nviews][numleafs] first reserved slot of a visited leaf, -1 = not visited */
	int32_t   *leafkey;                 /* [nviews][numleafs] place of a visited leaf in the walk (translucent brush entities) */
	/* translucent brush entities (k_tbrush_pieces -> k_tris_ordered): per (view, entity) an edge frame of its own */
	const int32_t *tbord; int32_t ntb;  /* bent index -> ordinal among the translucent ones (-1: opaque) */
	rc_surf_t *tb_surfs; rc_edge_t *tb_edges; int32_t *tb_counts; /* [nviews][ntb][RC_TB_MAXSURFS | RC_TB_MAXEDGES | 2 = pieces, edges] */
	int32_t   *viewbm;                  /* [nviews] k_bmodels posted a piece in the view: its scanlines check their events for pieces */
	uint32_t  *bmvis;                   /* [nviews][nodewords + leafwords] node / leaf visframe bits of the view */
	int32_t   *partbox;                 /* [nparts][6] bounding box of the part's transformed vertices as ordered int keys */
	uint8_t   *itemvis;                 /* [nviews][nitems] k_item_cull: 0 = no triangle of the item can be drawn in the view */
};

#define SP_LINES 32 /* scanlines per band (rc_k_spans.cuh): k_faces bins edges by band */
#define LN_GRID  ( 148 * 12 ) /* k_lines blocks: 12 resident blocks of 4 warps per SM */
#define RC_MAX_SLICES 4 /* view slices of a batch that run on concurrent streams */
#define CACHE_THREADS 256
#define CACHE_LUXCAP 2048

__device__ __forceinline__ int32_t rc_warp_incl_scan( int32_t x, int lane )
{
	for( int s = 1; s < 32; s <<= 1 )
	{
		int32_t t = __shfl_up_sync( 0xFFFFFFFFu, x, s );
		if( lane >= s ) x += t;
	}
	return x;
}

/* ---------------------------------------------------------------------------------------------- */
/* S0: visibility + BSP walk                                                                         */

#include "rc_bulk.cuh"
#include "rc_k_vis.cuh"

/* ---------------------------------------------------------------------------------------------- */
/* S1: faces -> surf_t records + edges                                                               */

#define FC_THREADS 64

/* the end of a face: side passes, surf_t record, unused edge reservations, the cache block it samples, counters */
__device__ __forceinline__ void rc_face_commit( const rc_scene_t &sc, const rc_work_t &w, int view, int ci, const rc_cand_t *cand,
	rc_facestate_t *fs, int ml, int mr, rc_surf_t *out )
{
	(void)ci;
	rc_surf_t o;
	o.miplevel = 0; o.cache_ofs = -1; o.cachewidth = 0;
	o.d_zistepu = o.d_zistepv = o.d_ziorigin = 0;
	o.d_sdivzstepu = o.d_tdivzstepu = o.d_sdivzstepv = o.d_tdivzstepv = o.d_sdivzorigin = o.d_tdivzorigin = 0;
	o.sadjust = o.tadjust = o.bbextents = o.bbextentt = 0;
	rc_face_finish( &sc, &w.views[view], cand, fs, ml, mr, &o );
	*out = o;

	int32_t nout = fs->nout < fs->maxout ? fs->nout : fs->maxout;
	/* unused reservations become edges that are never active */
	for( int32_t k = nout; k < fs->maxout; k++ )
	{
		fs->out[k].u = 0; fs->out[k].u_step = 0; fs->out[k].vv = 1u; fs->out[k].surfs = 0;
	}
	if( o.emitted && !( o.flags & ( RC_SURF_DRAWSKY | RC_SURF_DRAWTURB )) && o.cache_ofs >= 0 )
	{
		int32_t slot = cand->surf * 4 + o.miplevel;
		atomicOr( &w.cache_need[slot >> 5], 1u << ( slot & 31 ));
	}
	/* batch counters: one atomic per warp (tens of thousands of threads adding to one address serialise in L2) */
	{
		const unsigned m = __activemask();
		const unsigned ns = __reduce_add_sync( m, o.emitted ? 1u : 0u ), ne = __reduce_add_sync( m, (unsigned)nout );
		if(( threadIdx.x & 31 ) == __ffs( (int)m ) - 1 )
		{
			if( ns ) atomicAdd( &w.counters[RC_CTR_SURFS], (unsigned long long)ns );
			if( ne ) atomicAdd( &w.counters[RC_CTR_EDGES], (unsigned long long)ne );
		}
	}
}

__global__ void __launch_bounds__( FC_THREADS ) k_faces( rc_scene_t sc, rc_work_t w )
{
	const int view = blockIdx.x;
	const int ci = blockIdx.y * blockDim.x + threadIdx.x;
	if( ci >= w.ncands[view] )
		return;
	const rc_view_t *vw = &w.views[view];
	const rc_cand_t cand = w.cands[(size_t)view * w.max_cands + ci];
	rc_surf_t *out = &w.surfs[(size_t)view * w.max_cands + ci];
	if( cand.surf < 0 )
	{
		/* a slot k_viswalk reserved for a brush entity's face in this leaf: k_bmodels fills it when a piece lands there */
		out->emitted = 0; out->flags = 0; out->surf = -1;
		w.facerecs[(size_t)view * w.max_cands + ci].wrote = 0; w.facerecs[(size_t)view * w.max_cands + ci].state = 0;
		w.facewrote[(size_t)view * w.max_cands + ci] = 0;
		return;
	}
	const rc_surface_t *sf = &sc.surfaces[cand.surf];

	/* reserve the most this face can emit: one edge per surfedge + the left re-entry edge */
	int32_t want = sf->numedges + 1;
	int32_t first = atomicAdd( &w.nedges[view], want );
	rc_facestate_t fs;
	fs.out = w.edges + (size_t)view * w.max_edges + first;
	fs.maxout = want; fs.nout = 0;
	fs.surfnum = (uint32_t)ci + 1;
	if( first + want > w.max_edges )
	{
		/* the part of the reservation that is inside the list must not stay uninitialised: k_events reads all of it */
		for( int32_t k = first; k < w.max_edges; k++ )
		{
			rc_edge_t *ne = w.edges + (size_t)view * w.max_edges + k;
			ne->u = 0; ne->u_step = 0; ne->vv = 1u; ne->surfs = 0;
		}
		fs.maxout = 0;
		atomicAdd( &w.counters[RC_CTR_OVERFLOW_EDGES], 1ull );
	}

	const rc_cand_t *vcands = w.cands + (size_t)view * w.max_cands;
	rc_facerec_t *recs = w.facerecs + (size_t)view * w.max_cands;
	int ml, mr;
	const int need = rc_face_begin( &sc, vw, vcands, w.candof + (size_t)view * ( sc.numsurfaces + 2 ), ci, &fs, &ml, &mr );
	if( need )
	{
		/* one of the clip points this face builds its left re-entry edge / right nearzi from is an earlier face's: the
		 * records of the faces before it are complete only when this kernel is: k_faces_fix finishes the face */
		const int32_t slot = atomicAdd( w.fix_count, 1 );
		if( slot < w.fix_cap )
		{
			rc_face_store( &fs, ml, mr, 1, &recs[ci] );
			w.facewrote[(size_t)view * w.max_cands + ci] = (uint8_t)fs.wrote;
			w.fix_list[3 * slot] = view; w.fix_list[3 * slot + 1] = ci; w.fix_list[3 * slot + 2] = first;
			return;
		}
		atomicAdd( &w.counters[RC_CTR_OVERFLOW_FIX], 1ull ); /* finished here with the point at the origin; reported */
	}
	rc_face_commit( sc, w, view, ci, &cand, &fs, ml, mr, out );
	rc_face_store( &fs, ml, mr, 0, &recs[ci] );
	w.facewrote[(size_t)view * w.max_cands + ci] = (uint8_t)fs.wrote;
}

/* opaque brush entities (rc_core.h A13): thread per (view, entity face).  Runs after k_faces / k_faces_fix: the pieces go
 * into the slots k_viswalk reserved at their leaf's place in the surface order, their edges behind the world's */
__global__ void __launch_bounds__( 64 ) k_bmodels( rc_scene_t sc, rc_work_t w )
{
	const int view = blockIdx.y;
	const int32_t g = blockIdx.x * blockDim.x + threadIdx.x;
	if( g >= w.nbfaces )
		return;
	const int32_t ent = w.bfaces[g] >> 16, lf = w.bfaces[g] & 0xFFFF;
	const rc_bent_t *be = &w.bents[ent];
	const rc_view_t *vw = &w.views[view];
	const int32_t clipflags = rc_bmodel_check_bbox( vw, be->minmaxs );
	if( clipflags < 0 )
		return;
	const uint32_t *nodevis = w.bmvis + (size_t)view * ( w.nodewords + w.leafwords ), *leafvis = nodevis + w.nodewords;
	const int32_t topnode = rc_find_topnode( &sc, nodevis, leafvis, be->minmaxs );
	if( topnode == RC_TOPNODE_NONE )
		return;
	rc_view_t bv;
	rc_bmodel_view( vw, be, &bv );
	rc_bmout_t out;
	out.surfs = w.surfs + (size_t)view * w.max_cands; out.max_cands = w.ncands[view];
	out.edges = w.edges + (size_t)view * w.max_edges; out.max_edges = w.max_edges; out.nedges = &w.nedges[view];
	out.leafbase = w.leafbase + (size_t)view * sc.numleafs;
	out.candof = w.candof + (size_t)view * ( sc.numsurfaces + 2 );
	out.cache_need = w.cache_need; out.leafkey = nullptr;
	out.nsurfs = 0; out.overflow = 0; out.noutmax = 0;
	rc_bmodel_face( &sc, &bv, w.bents, w.nbents, ent, be->firstface + lf, clipflags, topnode, nodevis, leafvis, &out );
	if( out.nsurfs )
	{
		w.viewbm[view] = 1;
		atomicAdd( &w.counters[RC_CTR_BMODEL_SURFS], (unsigned long long)out.nsurfs );
		atomicAdd( &w.counters[RC_CTR_SURFS], (unsigned long long)out.nsurfs );
	}
	if( out.overflow ) atomicAdd( &w.counters[RC_CTR_OVERFLOW_BMODEL], 1ull );
	if( out.noutmax > w.max_edges ) atomicAdd( &w.counters[RC_CTR_OVERFLOW_EDGES], 1ull );
}

/* translucent brush entities (rc_core.h A12): thread per (view, entity) - the pieces of R_DrawBrushModel's own edge frame,
 * kept per (view, entity) for k_tris_ordered, which draws them in the entity's turn; the mip-0 cache blocks they sample
 * join the batch's build list (this runs before k_cache_list) */
__global__ void __launch_bounds__( 64 ) k_tbrush_pieces( rc_scene_t sc, rc_work_t w )
{
	const int32_t g = blockIdx.x * blockDim.x + threadIdx.x;
	if( g >= w.nviews * w.nbents )
		return;
	const int view = g / w.nbents, ent = g - view * w.nbents;
	const int32_t tb = w.tbord[ent];
	if( tb < 0 )
		return;
	const size_t slot = (size_t)view * w.ntb + tb;
	const uint32_t *nodevis = w.bmvis + (size_t)view * ( w.nodewords + w.leafwords ), *leafvis = nodevis + w.nodewords;
	int32_t ovf = 0;
	int32_t *pnedges = &w.tb_counts[2 * slot + 1]; /* the edge cursor is advanced with an atomic: it has to live in global memory */
	const int32_t ns = rc_tbrush_pieces( &sc, &w.views[view], w.bents, w.nbents, ent, nodevis, leafvis, w.leafbase + (size_t)view * sc.numleafs,
		w.leafkey + (size_t)view * sc.numleafs, w.candof + (size_t)view * ( sc.numsurfaces + 2 ),
		w.tb_surfs + slot * RC_TB_MAXSURFS, w.tb_edges + slot * RC_TB_MAXEDGES, pnedges, w.cache_need, &ovf );
	w.tb_counts[2 * slot] = ns;
	if( ns ) atomicAdd( &w.counters[RC_CTR_BMODEL_SURFS], (unsigned long long)ns );
	if( ovf ) atomicAdd( &w.counters[RC_CTR_OVERFLOW_BMODEL], 1ull );
}

/* the faces k_faces left waiting for an earlier face's clip point (a few per view): a warp per face - the lanes look at
 * the 32 faces before it at a time (one byte each: which of the four points that face wrote), the nearest writer of
 * every missing point is picked with a ballot; lane 0 then finishes the face */
__global__ void __launch_bounds__( FC_THREADS ) k_faces_fix( rc_scene_t sc, rc_work_t w )
{
	int32_t n = *w.fix_count;
	if( n > w.fix_cap ) n = w.fix_cap;
	const int lane = threadIdx.x & 31;
	const int32_t nwarps = gridDim.x * ( FC_THREADS / 32 );
	for( int32_t i = blockIdx.x * ( FC_THREADS / 32 ) + ( threadIdx.x >> 5 ); i < n; i += nwarps )
	{
		const int32_t view = w.fix_list[3 * i], ci = w.fix_list[3 * i + 1], first = w.fix_list[3 * i + 2];
		const rc_facerec_t *recs = w.facerecs + (size_t)view * w.max_cands;
		const uint8_t *wrote = w.facewrote + (size_t)view * w.max_cands;
		const rc_facerec_t me = recs[ci];
		const int ml = me.state & 1, mr = ( me.state >> 1 ) & 1;
		int need = rc_face_need( ml, mr, me.wrote );
		float pts[4][3];
		for( int b = 0; b < 4; b++ ) for( int q = 0; q < 3; q++ ) pts[b][q] = me.pts[b][q];
		for( int32_t base = ci - 1; base >= 0 && need; base -= 32 )
		{
			const int32_t k = base - lane;
			const int take = k >= 0 ? ( (int)wrote[k] & need ) : 0;
#pragma unroll
			for( int b = 0; b < 4; b++ )
			{
				const uint32_t m = __ballot_sync( 0xFFFFFFFFu, take & ( 1 << b ));
				if( m )
				{
					const rc_facerec_t *src = &recs[base - ( __ffs( (int)m ) - 1 )]; /* the lowest lane is the nearest face */
					for( int q = 0; q < 3; q++ ) pts[b][q] = src->pts[b][q];
					need &= ~( 1 << b );
				}
			}
		}
		if( lane == 0 )
		{
			const rc_cand_t cand = w.cands[(size_t)view * w.max_cands + ci];
			const rc_surface_t *sf = &sc.surfaces[cand.surf];
			rc_facestate_t fs;
			fs.out = w.edges + (size_t)view * w.max_edges + first;
			fs.maxout = first + sf->numedges + 1 > w.max_edges ? 0 : sf->numedges + 1;
			fs.nout = me.nout;
			fs.surfnum = (uint32_t)ci + 1;
			fs.emitted = ( me.state & 4 ) ? 1 : 0; fs.nearzi = me.nearzi; fs.nearzionly = 0; fs.lastvertvalid = 0; fs.wrote = me.wrote;
			fs.u1 = fs.v1 = fs.lzi1 = 0; fs.ceilv1 = 0; fs.leftclipped = fs.rightclipped = 0;
			for( int q = 0; q < 3; q++ )
			{
				fs.leftexit[q] = pts[0][q]; fs.leftenter[q] = pts[1][q]; fs.rightexit[q] = pts[2][q]; fs.rightenter[q] = pts[3][q];
			}
			rc_face_commit( sc, w, view, ci, &cand, &fs, ml, mr, &w.surfs[(size_t)view * w.max_cands + ci] );
			atomicAdd( &w.counters[RC_CTR_FIXED_FACES], 1ull );
		}
		__syncwarp();
	}
}

/* ---------------------------------------------------------------------------------------------- */
/* S2: surface cache                                                                                 */

__device__ __forceinline__ bool rc_key_stale( const rc_cachekey_t *k, const rc_surface_t *sf, const int32_t *styles, uint32_t decalgen )
{
	if( !k->valid || k->texnum != sf->texnum || (uint32_t)k->decalgen != decalgen )
		return true;
	for( int j = 0; j < RC_MAXLIGHTMAPS && sf->styles[j] != 255; j++ )
		if( k->light[j] != styles[sf->styles[j]] )
			return true;
	return false;
}

#define CACHE_FLUX 256 /* luxels / light blocks of an item the flat path takes (a 16 x 16 lightmap: faces up to 240 texels) */

/* geometry of a (face, mip) block; fast = made of whole light blocks (every face whose extents are multiples of 16) */
struct rc_cacheitem_t { int32_t smax, tmax, size, bw, bh, nbu, nbv, rl, nruns; bool fast; };
__device__ __forceinline__ rc_cacheitem_t rc_cache_item( const rc_surface_t *sf, int32_t mip )
{
	rc_cacheitem_t c;
	c.smax = sf->lightextents[0] / sf->sample_size + 1; c.tmax = sf->lightextents[1] / sf->sample_size + 1;
	c.size = c.smax * c.tmax;
	c.bw = sf->lightextents[0] >> mip; c.bh = sf->lightextents[1] >> mip;
	const int32_t k = 4 - mip, bs = 16 >> mip;
	c.nbu = c.bw >> k; c.nbv = c.bh >> k;
	c.rl = mip <= 1 ? 8 : ( mip == 2 ? 4 : 2 );
	c.nruns = ( c.bw / c.rl ) * c.bh;
	c.fast = c.nbu * bs == c.bw && c.nbv * bs == c.bh && c.nbu < c.smax && c.nbv < c.tmax && c.size <= CACHE_FLUX && c.nbu * c.nbv <= CACHE_FLUX
		&& c.nruns > 0 && c.nruns <= 32 * 256;
	return c;
}

__global__ void __launch_bounds__( 256 ) k_cache_list( rc_scene_t sc, rc_work_t w )
{
	const int slot = blockIdx.x * blockDim.x + threadIdx.x;
	if( slot >= sc.numsurfaces * 4 )
		return;
	if( !(( w.cache_need[slot >> 5] >> ( slot & 31 )) & 1 ))
		return;
	const rc_surface_t *sf = &sc.surfaces[slot >> 2];
	if( sf->cache_ofs[slot & 3] < 0 )
		return;
	/* a face lit by a dynamic light is rebuilt every frame, and so is a block that was built with one */
	const uint32_t decalgen = w.decal_gen ? w.decal_gen[slot >> 2] : 0u;
	if( !( rc_key_stale( &w.cache_keys[slot], sf, w.styles, decalgen ) || w.cache_keys[slot].dlight || ( w.ndlights && w.dlightbits[slot >> 2] )))
		return;
	const rc_cacheitem_t c = rc_cache_item( sf, slot & 3 );
	if( !c.fast || ( w.decal_first && w.decal_first[2 * ( slot >> 2 ) + 1] > 0 ))
	{
		/* faces that carry decals take the block-per-item path too: it blits them after the texels */
		w.cache_list[atomicAdd( w.cache_nlist, 1 )] = slot; /* ragged blocks: k_cache_build, a CTA per item */
		return;
	}
	const int32_t pos = atomicAdd( &w.cache_fn[0], 1 );
	w.cache_flist[pos] = slot;
	const int32_t npieces = ( c.nruns + 255 ) >> 8;
	const int32_t u0 = atomicAdd( &w.cache_fn[1], npieces );
	for( int32_t u = 0; u < npieces; u++ ) w.cache_units[u0 + u] = ((uint32_t)pos << 6 ) | (uint32_t)u;
}

/* blocklights and light-block source offsets of every listed item: a warp per item, lanes over the luxels / blocks */
__global__ void __launch_bounds__( 256 ) k_cache_light( rc_scene_t sc, rc_work_t w )
{
	const int lane = threadIdx.x & 31;
	const int32_t n = w.cache_fn[0], nwarps = gridDim.x * ( blockDim.x >> 5 );
	for( int32_t it = blockIdx.x * ( blockDim.x >> 5 ) + ( threadIdx.x >> 5 ); it < n; it += nwarps )
	{
		const int32_t slot = w.cache_flist[it], mip = slot & 3;
		const rc_surface_t *sf = &sc.surfaces[slot >> 2];
		const rc_texture_t *tex = &sc.textures[sf->texnum];
		const rc_cacheitem_t c = rc_cache_item( sf, mip );
		const uint32_t dbits = w.ndlights ? w.dlightbits[slot >> 2] : 0u;
		uint32_t *lux = w.cache_lux + (size_t)slot * CACHE_FLUX;
		for( int32_t i = lane; i < c.size; i += 32 )
		{
			uint32_t extra = 0;
			if( dbits )
			{
				const int32_t lt = i / c.smax;
				extra = rc_dlight_luxel( &sc, sf, w.dlights, w.ndlights, dbits, i - lt * c.smax, lt );
			}
			lux[i] = rc_blocklight( &sc, sf, w.styles, c.size, i, extra );
		}
		/* source offset of every light block: the column offset advances per block column and snaps to 0 past the texture
		 * width, the row offset advances per block row and steps back by the texture size (r_surf.c:411-426,586-600) */
		const int32_t bs = 16 >> mip;
		int32_t smaxt = tex->width >> mip, tmaxt = tex->height >> mip;
		if( smaxt < 1 ) smaxt = 1;
		if( tmaxt < 1 ) tmaxt = 1;
		const int32_t total = tmaxt * smaxt;
		int32_t *srcofs = w.cache_src + (size_t)slot * CACHE_FLUX;
		for( int32_t bu = lane; bu < c.nbu; bu += 32 )
		{
			int32_t soffset = ((( (int32_t)sf->lightmapmins[0] ) >> mip ) + ( smaxt << 16 )) % smaxt;
			for( int32_t q = 0; q < bu; q++ )
			{
				soffset += bs;
				if( soffset >= smaxt ) soffset = 0;
			}
			int32_t rowofs = (((( (int32_t)sf->lightmapmins[1] ) >> mip ) + ( tmaxt << 16 )) % tmaxt ) * smaxt;
			for( int32_t bv = 0; bv < c.nbv; bv++ )
			{
				srcofs[bv * c.nbu + bu] = rowofs + soffset;
				rowofs += bs * smaxt;
				if( rowofs + soffset >= total ) rowofs -= total;
			}
		}
	}
}

/*
 * R_DrawSurface + R_DrawSurfaceBlock8_mipN (r_surf.c:291-757) for runs of RL texels of one row inside one light
 * block: the light ramp of the row is set up once per run (uint32 wraparound arithmetic as in the reference), the RL
 * texels leave as one 4/8/16-byte store.  s_rc[bv * nbu + bu] = source offset of the block's first texel (the
 * reference's psource after its column / row wrap rules, rc_cache_texel).
 */
template <int RL>
__device__ __forceinline__ void rc_cache_run( const uint16_t *__restrict__ src, const uint16_t *__restrict__ cmap, int32_t lim, int32_t smaxt, int32_t mip,
	const uint32_t *bl, int32_t lightwidth, const int32_t *rc, int32_t nbu, uint16_t *dst, int32_t bw, int32_t rpr, int32_t r )
{
	const int32_t k = 4 - mip, bs = 16 >> mip;
	const int32_t y = r / rpr, x = ( r - y * rpr ) * RL;
	const int32_t bu = x >> k, b0 = x & ( bs - 1 ), bv = y >> k, i = y & ( bs - 1 );
	const uint32_t *lp = bl + bv * lightwidth + bu;
	uint32_t ll = lp[0], lr = lp[1];
	const uint32_t lls = ( lp[lightwidth] - ll ) >> k, lrs = ( lp[lightwidth + 1] - lr ) >> k;
	ll += (uint32_t)i * lls; lr += (uint32_t)i * lrs;
	const uint32_t step = ( ll - lr ) >> k;
	uint32_t light = lr + (uint32_t)( bs - 1 - b0 ) * step;
	const int32_t idx0 = rc[bv * nbu + bu] + i * smaxt + b0;
	uint32_t packed[RL / 2];
#pragma unroll
	for( int j = 0; j < RL; j++ )
	{
		/* the reference reads past the mip when a texture is not a multiple of the block size; stay inside the arena */
		int32_t idx = idx0 + j;
		idx = idx < 0 ? 0 : ( idx >= lim ? lim - 1 : idx );
		const uint32_t pix = __ldg( src + idx );
		uint32_t o = (uint32_t)__ldg( cmap + (( pix >> 3 ) | (( light & 0x1f00u ) << 5 ))) | ( pix & 7u );
		if( mip == 0 && pix == RC_TRANSPARENT_COLOR )
			o = RC_TRANSPARENT_COLOR;
		if( j & 1 ) packed[j >> 1] |= o << 16; else packed[j >> 1] = o;
		light -= step;
	}
	uint16_t *out = dst + (size_t)y * bw + x;
	if( RL == 8 ) *reinterpret_cast<uint4 *>( out ) = make_uint4( packed[0], packed[1], packed[2 % ( RL / 2 )], packed[3 % ( RL / 2 )] );
	else if( RL == 4 ) *reinterpret_cast<uint2 *>( out ) = make_uint2( packed[0], packed[1 % ( RL / 2 )] );
	else *reinterpret_cast<uint32_t *>( out ) = packed[0];
}

template <int RL>
__device__ __forceinline__ void rc_cache_runs( const rc_scene_t &sc, const rc_texture_t *tex, int32_t mip, const uint32_t *s_bl, int32_t lightwidth,
	const int32_t *s_rc, int32_t nbu, uint16_t *dst, int32_t bw, int32_t bh, int tid )
{
	int32_t smaxt = tex->width >> mip;
	if( smaxt < 1 ) smaxt = 1;
	const int32_t lim = tex->width * tex->height * 2;
	const uint16_t *__restrict__ src = sc.texels + tex->ofs[mip];
	const int32_t rpr = bw / RL, total = rpr * bh;
	for( int32_t r = tid; r < total; r += CACHE_THREADS )
		rc_cache_run<RL>( src, sc.colormap, lim, smaxt, mip, s_bl, lightwidth, s_rc, nbu, dst, bw, rpr, r );
}

/* the texels of the listed items, flat: a CTA per piece of 256 runs, a thread per run of 8 / 4 / 2 texels of one row of one
 * light block (one 16 / 8 / 4-byte store); blocklights and source offsets come from k_cache_light's tables (L1 / L2) */
__global__ void __launch_bounds__( 256 ) k_cache_texels( rc_scene_t sc, rc_work_t w )
{
	const int32_t npieces = w.cache_fn[1];
	for( int32_t u = blockIdx.x; u < npieces; u += gridDim.x )
	{
		const uint32_t e = w.cache_units[u];
		const int32_t slot = w.cache_flist[e >> 6], mip = slot & 3;
		const rc_surface_t *sf = &sc.surfaces[slot >> 2];
		const rc_texture_t *tex = &sc.textures[sf->texnum];
		const int32_t smax = sf->lightextents[0] / sf->sample_size + 1;
		const int32_t bw = sf->lightextents[0] >> mip, bh = sf->lightextents[1] >> mip, nbu = bw >> ( 4 - mip );
		int32_t smaxt = tex->width >> mip;
		if( smaxt < 1 ) smaxt = 1;
		const int32_t lim = tex->width * tex->height * 2;
		const uint16_t *src = sc.texels + tex->ofs[mip];
		uint16_t *dst = sc.cache + sf->cache_ofs[mip];
		const uint32_t *bl = w.cache_lux + (size_t)slot * CACHE_FLUX;
		const int32_t *rc = w.cache_src + (size_t)slot * CACHE_FLUX;
		const int32_t r = (int32_t)( e & 63u ) * 256 + threadIdx.x;
		if( mip <= 1 ) { const int32_t rpr = bw / 8; if( r < rpr * bh ) rc_cache_run<8>( src, sc.colormap, lim, smaxt, mip, bl, smax, rc, nbu, dst, bw, rpr, r ); }
		else if( mip == 2 ) { const int32_t rpr = bw / 4; if( r < rpr * bh ) rc_cache_run<4>( src, sc.colormap, lim, smaxt, mip, bl, smax, rc, nbu, dst, bw, rpr, r ); }
		else { const int32_t rpr = bw / 2; if( r < rpr * bh ) rc_cache_run<2>( src, sc.colormap, lim, smaxt, mip, bl, smax, rc, nbu, dst, bw, rpr, r ); }
	}
}

/* the keys of the blocks k_cache_texels wrote (after it, in stream order: a key says "these texels are current") */
__global__ void __launch_bounds__( 256 ) k_cache_publish( rc_scene_t sc, rc_work_t w )
{
	const int32_t it = blockIdx.x * blockDim.x + threadIdx.x;
	unsigned long long texels = 0;
	if( it < w.cache_fn[0] )
	{
		const int32_t slot = w.cache_flist[it], mip = slot & 3;
		const rc_surface_t *sf = &sc.surfaces[slot >> 2];
		rc_cachekey_t *kk = &w.cache_keys[slot];
		kk->texnum = sf->texnum;
		for( int j = 0; j < RC_MAXLIGHTMAPS; j++ )
			kk->light[j] = sf->styles[j] != 255 ? w.styles[sf->styles[j]] : 0;
		kk->dlight = ( w.ndlights && w.dlightbits[slot >> 2] ) ? 1 : 0;
		kk->decalgen = (int32_t)( w.decal_gen ? w.decal_gen[slot >> 2] : 0u );
		kk->valid = 1;
		texels = (unsigned long long)( sf->lightextents[0] >> mip ) * ( sf->lightextents[1] >> mip );
	}
	/* batch counters: one atomic pair per warp */
	const unsigned m = __activemask();
	const unsigned long long tsum = __reduce_add_sync( m, (unsigned)texels );
	const unsigned nb = __popc( __ballot_sync( m, texels != 0 ));
	if(( threadIdx.x & 31 ) == __ffs( (int)m ) - 1 && nb )
	{
		atomicAdd( &w.counters[RC_CTR_CACHE_BLOCKS], (unsigned long long)nb );
		atomicAdd( &w.counters[RC_CTR_CACHE_TEXELS], tsum );
	}
}

__global__ void __launch_bounds__( CACHE_THREADS ) k_cache_build( rc_scene_t sc, rc_work_t w )
{
	__shared__ uint32_t s_bl[CACHE_LUXCAP];
	__shared__ int32_t s_rc[CACHE_LUXCAP];
	__shared__ int32_t s_item;
	const int tid = threadIdx.x;
	const int32_t nlist = *w.cache_nlist;

	for( ;; )
	{
		__syncthreads();
		if( tid == 0 ) s_item = atomicAdd( w.cache_cursor, 1 );
		__syncthreads();
		const int32_t item = s_item;
		if( item >= nlist )
			break;
		const int32_t slot = w.cache_list[item];
		const int32_t mip = slot & 3;
		const rc_surface_t *sf = &sc.surfaces[slot >> 2];
		const rc_texture_t *tex = &sc.textures[sf->texnum];
		const int32_t smax = sf->lightextents[0] / sf->sample_size + 1, tmax = sf->lightextents[1] / sf->sample_size + 1;
		const int32_t size = smax * tmax;
		if( size > CACHE_LUXCAP )
		{
			if( tid == 0 ) atomicAdd( &w.counters[RC_CTR_OVERFLOW_CACHE], 1ull );
			continue;
		}
		const uint32_t dbits = w.ndlights ? w.dlightbits[slot >> 2] : 0u;
		for( int32_t i = tid; i < size; i += CACHE_THREADS )
		{
			uint32_t extra = 0;
			if( dbits )
			{
				const int32_t lt = i / smax;
				extra = rc_dlight_luxel( &sc, sf, w.dlights, w.ndlights, dbits, i - lt * smax, lt );
			}
			s_bl[i] = rc_blocklight( &sc, sf, w.styles, size, i, extra );
		}

		const int32_t bw = sf->lightextents[0] >> mip, bh = sf->lightextents[1] >> mip;
		const int32_t k = 4 - mip, bs = 16 >> mip;
		const int32_t nbu = bw >> k, nbv = bh >> k;
		uint16_t *dst = sc.cache + sf->cache_ofs[mip];
		const bool fast = nbu * bs == bw && nbv * bs == bh && nbu < smax && nbv < tmax && nbu * nbv <= CACHE_LUXCAP;
		if( fast )
		{
			/* source offset of every light block: column offset advances per block column and snaps to 0 past the
			 * texture width, row offset advances per block row and steps back by the texture size (r_surf.c:411-426,586-600) */
			int32_t smaxt = tex->width >> mip, tmaxt = tex->height >> mip;
			if( smaxt < 1 ) smaxt = 1;
			if( tmaxt < 1 ) tmaxt = 1;
			const int32_t total = tmaxt * smaxt;
			for( int32_t bu = tid; bu < nbu; bu += CACHE_THREADS )
			{
				int32_t soffset = ((( (int32_t)sf->lightmapmins[0] ) >> mip ) + ( smaxt << 16 )) % smaxt;
				for( int32_t q = 0; q < bu; q++ )
				{
					soffset += bs;
					if( soffset >= smaxt ) soffset = 0;
				}
				int32_t rowofs = (((( (int32_t)sf->lightmapmins[1] ) >> mip ) + ( tmaxt << 16 )) % tmaxt ) * smaxt;
				for( int32_t bv = 0; bv < nbv; bv++ )
				{
					s_rc[bv * nbu + bu] = rowofs + soffset;
					rowofs += bs * smaxt;
					if( rowofs + soffset >= total ) rowofs -= total;
				}
			}
		}
		__syncthreads();

		/* R_DrawSurfaceDecals (r_surf.c:914): the face's decals over the lit texel, in list order, before the one store
		 * of the texel (another slice may be building the same block at the same time: nothing is read back from it) */
		const int32_t ndecals = w.decal_first ? w.decal_first[2 * ( slot >> 2 ) + 1] : 0;
		if( !fast || ndecals > 0 )
		{
			const rc_ddecal_t *dl = ndecals > 0 ? w.decals + w.decal_first[2 * ( slot >> 2 )] : nullptr;
			rc_blendluts_t L;
			L.colormap = sc.colormap; L.alphamap = w.alphamap; L.addmap = w.addmap; L.modmap = w.modmap;
			for( int32_t i = tid; i < bw * bh; i += CACHE_THREADS )
			{
				int32_t y = i / bw, x = i - y * bw;
				uint32_t px = rc_cache_texel( &sc, sf, tex, mip, s_bl, smax, x, y );
				if( ndecals > 0 )
					rc_decal_texel( &L, sc.texels, dl, ndecals, mip, x, y, &px );
				dst[i] = (uint16_t)px;
			}
		}
		else if( mip <= 1 ) rc_cache_runs<8>( sc, tex, mip, s_bl, smax, s_rc, nbu, dst, bw, bh, tid );
		else if( mip == 2 ) rc_cache_runs<4>( sc, tex, mip, s_bl, smax, s_rc, nbu, dst, bw, bh, tid );
		else rc_cache_runs<2>( sc, tex, mip, s_bl, smax, s_rc, nbu, dst, bw, bh, tid );

		/* another slice's stream may test this key while the block is being written: publish it only after the texels */
		__threadfence();
		__syncthreads();
		if( tid == 0 )
		{
			rc_cachekey_t *kk = &w.cache_keys[slot];
			kk->texnum = sf->texnum;
			for( int j = 0; j < RC_MAXLIGHTMAPS; j++ )
				kk->light[j] = sf->styles[j] != 255 ? w.styles[sf->styles[j]] : 0;
			kk->decalgen = (int32_t)( w.decal_gen ? w.decal_gen[slot >> 2] : 0u );
			kk->valid = 1;
			kk->dlight = dbits ? 1 : 0;
			atomicAdd( &w.counters[RC_CTR_CACHE_BLOCKS], 1ull );
			atomicAdd( &w.counters[RC_CTR_CACHE_TEXELS], (unsigned long long)bw * bh );
		}
	}
}

#include "rc_k_spans.cuh"
#include "rc_k_draw.cuh"
#include "rc_k_tris.cuh"

/* ---------------------------------------------------------------------------------------------- */
/* host side of the C layer                                                                          */

namespace {

/* ---------------------------------------------------------------------------------------------- */
/* R_PushDlights / R_MarkLights (ref/common/ref_light.c:93-185) for the world: thread per surface.  The recursion
 * reaches the node a surface hangs off iff at every ancestor the light's distance to the split plane does not send
 * it down the other side only; at the node itself the reference's plane-distance and clamped-impact tests apply.    */

__global__ void __launch_bounds__( 128 ) k_dlight_mark( rc_scene_t sc, const rc_dlight_t *lights, int32_t nlights, uint32_t *bits )
{
	const int s = blockIdx.x * blockDim.x + threadIdx.x;
	if( s >= sc.numsurfaces )
		return;
	const rc_surface_t *sf = &sc.surfaces[s];
	const int32_t n0 = sc.surfnode[s];
	uint32_t out = 0;
	/* a surface no node lists is never marked (surfnode defaults to 0 for those: check the range) */
	const bool listed = s >= sc.nodes[n0].firstsurface && s < sc.nodes[n0].firstsurface + sc.nodes[n0].numsurfaces;
	for( int32_t k = 0; k < nlights && listed; k++ )
	{
		const rc_dlight_t *dl = &lights[k];
		const float vr = dl->virtual_radius, maxdist = dl->radius * dl->radius;
		const float dist = rc_planediff( dl->origin, &sc.planes[sc.nodes[n0].plane] );
		if( dist > vr || dist < -vr )
			continue; /* the walk passes this node without looking at its surfaces */
		const float dist_sq = dist * dist;
		if( !( dist_sq < maxdist ))
			continue;
		bool reach = true;
		for( int32_t c = n0, a = sc.nodes[n0].parent; a >= 0 && reach; c = a, a = sc.nodes[a].parent )
		{
			const float d = rc_planediff( dl->origin, &sc.planes[sc.nodes[a].plane] );
			if( d > vr ) reach = sc.nodes[a].child[0] == c;
			else if( d < -vr ) reach = sc.nodes[a].child[1] == c;
		}
		if( !reach )
			continue;
		const rc_plane_t *pl = &sc.planes[sf->plane];
		float impact[3];
		if( pl->type < 3 )
		{
			impact[0] = dl->origin[0]; impact[1] = dl->origin[1]; impact[2] = dl->origin[2];
			impact[pl->type] -= dist;
		}
		else
		{
			const float nd = -dist;
			impact[0] = dl->origin[0] + nd * pl->normal[0];
			impact[1] = dl->origin[1] + nd * pl->normal[1];
			impact[2] = dl->origin[2] + nd * pl->normal[2];
		}
		/* clamp the centre of the light to the face and check the brightness there */
		float l = rc_dot3( impact, sf->lmvecs[0] ) + sf->lmvecs[0][3] - (float)sf->lightmapmins[0];
		float sx = l + 0.5f;
		sx = sx >= 0 ? ( sx < (float)sf->lightextents[0] ? sx : (float)sf->lightextents[0] ) : 0;
		sx = l - sx;
		l = rc_dot3( impact, sf->lmvecs[1] ) + sf->lmvecs[1][3] - (float)sf->lightmapmins[1];
		float tx = l + 0.5f;
		tx = tx >= 0 ? ( tx < (float)sf->lightextents[1] ? tx : (float)sf->lightextents[1] ) : 0;
		tx = l - tx;
		if( sx * sx + tx * tx + dist_sq >= maxdist )
			continue;
		out |= 1u << dl->bit;
	}
	bits[s] = out;
}

/* ---------------------------------------------------------------------------------------------- */
/* display tail: R_BlitScreen (ref/soft/r_glblit.c:591-721) - every pixel_t goes through the 65536-entry
 * screen map (vid.screen / vid.screen32, built like R_BuildScreenMap by the host side).  HBM bound:
 * 2 B read + bpp B written per pixel; the table (256 KB) lives in L2.  Thread = 8 pixels.            */

template <int BPP>
__global__ void __launch_bounds__( 256 ) k_blit( const uint16_t *__restrict__ color, const uint32_t *__restrict__ lut, void *out, size_t npx )
{
	const size_t n8 = npx >> 3;
	for( size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (size_t)gridDim.x * blockDim.x )
	{
		const uint4 in = __ldg( reinterpret_cast<const uint4 *>( color ) + i );
		const uint32_t p[4] = { in.x, in.y, in.z, in.w };
		uint32_t o[8];
#pragma unroll
		for( int k = 0; k < 4; k++ )
		{
			o[2 * k] = __ldg( lut + ( p[k] & 0xFFFFu ));
			o[2 * k + 1] = __ldg( lut + ( p[k] >> 16 ));
		}
		if( BPP == 2 )
			reinterpret_cast<uint4 *>( out )[i] = make_uint4( o[0] | ( o[1] << 16 ), o[2] | ( o[3] << 16 ), o[4] | ( o[5] << 16 ), o[6] | ( o[7] << 16 ));
		else
		{
			reinterpret_cast<uint4 *>( out )[2 * i] = make_uint4( o[0], o[1], o[2], o[3] );
			reinterpret_cast<uint4 *>( out )[2 * i + 1] = make_uint4( o[4], o[5], o[6], o[7] );
		}
	}
	/* ragged tail (frame sizes that are not a multiple of 8 pixels) */
	if( blockIdx.x == 0 )
		for( size_t x = ( n8 << 3 ) + threadIdx.x; x < npx; x += blockDim.x )
		{
			const uint32_t v = __ldg( lut + color[x] );
			if( BPP == 2 ) reinterpret_cast<uint16_t *>( out )[x] = (uint16_t)v;
			else reinterpret_cast<uint32_t *>( out )[x] = v;
		}
}

/* fills of the batch path as kernels: a cudaMemsetAsync may be served by a copy engine, where it would queue behind the
 * frames another batch is copying to the host (measured: every stage of an asynchronous batch started 0.7 ms late) */
__global__ void k_fill32( uint32_t *dst, uint32_t value, size_t n )
{
	for( size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x ) dst[i] = value;
}
__global__ void k_copy32( uint32_t *dst, const uint32_t *src, size_t n )
{
	for( size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x ) dst[i] = src[i];
}
static inline void fill32( void *dst, uint32_t value, size_t nwords, cudaStream_t st )
{
	if( !nwords ) return;
	size_t blocks = ( nwords + 255 ) / 256;
	if( blocks > 148 * 8 ) blocks = 148 * 8;
	k_fill32<<<(unsigned)blocks, 256, 0, st>>>( (uint32_t *)dst, value, nwords );
}

/* the batch counters into their mapped pinned mirror: a store over PCIe instead of a copy that would queue behind the frames
 * on the device->host copy engine (the host reads the mirror once the batch's kernels are done, before its copies are) */
__global__ void k_counters_out( const unsigned long long *src, unsigned long long *dst )
{
	dst[threadIdx.x] = src[threadIdx.x];
	__threadfence_system();
}

struct Dev
{
	bool ready = false;
	int device = 0, num_sms = 148;
	char name[256] = "";
	char err[512] = "";
	cudaStream_t stream = nullptr, copy = nullptr;
	cudaStream_t slice_stream[RC_MAX_SLICES] = {};
	cudaEvent_t ev_slice[RC_MAX_SLICES] = {}, ev_fork = nullptr;
	cudaStream_t aux_stream[RC_MAX_SLICES] = {};           /* surface-cache kernels of a slice, beside its span kernels */
	cudaEvent_t ev_cache[RC_MAX_SLICES][2] = {};           /* start / end of those */
	bool cache_aside = true;                               /* false: on the slice's own stream (the bench times the kernels alone) */
	cudaEvent_t ev[10] = {};
	uint64_t launches = 0, h2d_bytes = 0;
	bool sync_each = false; /* RC_SYNC_EACH in the environment: synchronise and check after every kernel (names the one that faults) */

	rc_scene_t sc{};            /* device pointers */
	std::vector<rc_surface_t> h_surfaces;
	int32_t h_styles[256];
	size_t cache_texels = 0;
	int ntextures = 0;
	std::vector<rc_texture_t> h_textures;

	/* LUTs + static arrays */
	uint16_t *d_colormap = nullptr, *d_lightgamma = nullptr;
	int32_t *d_sintable = nullptr;
	rc_texture_t *d_textures = nullptr;
	uint16_t *d_texels = nullptr;
	void *d_scene[16] = {};
	int32_t *d_styles = nullptr;
	rc_cachekey_t *d_keys = nullptr;
	uint32_t *d_need = nullptr;
	int32_t *d_list = nullptr, *d_nlist = nullptr;
	int32_t *d_flist = nullptr, *d_cachesrc = nullptr; uint32_t *d_units = nullptr, *d_cachelux = nullptr;
	unsigned long long *d_counters = nullptr;

	/* batch work buffers (grown on demand) */
	rc_work_t w{};
	int cap_views = 0, cap_w = 0, cap_h = 0;
	/* asynchronous batches (rc_dev_render_async): a second set of frame buffers, so that batch k's frames can be copied out
	 * while batch k + 1 is drawn into the other set; the batch whose kernels have not been checked for overflows yet */
	uint16_t *main_color = nullptr, *alt_color = nullptr; int16_t *main_zbuf = nullptr, *alt_zbuf = nullptr; size_t alt_px = 0; bool use_alt = false;
	cudaEvent_t ev_kdone = nullptr; bool pending = false;
	struct batch_graph_t { cudaGraphExec_t exec = nullptr; unsigned long long key = 0, launches = 0; };
	batch_graph_t graphs[4][2]; /* [segment][frame buffer set] */
	bool use_graphs = true;
	bool vw_wide = false; /* k_viswalk<1024> for few views of a large tree */
	int ev_threads = 0; /* RC_EV_THREADS: k_events block size forced (measurement only) */
	int tr_strips = TR_STRIPS; /* horizontal strips of a frame in k_tris_ordered (RC_TR_STRIPS: measurement only) */
	cudaEvent_t ev_join[2] = {};
	std::vector<rc_view_t> pend_views; int pend_w = 0, pend_h = 0, pend_clear = 0; uint16_t *pend_color = nullptr; int16_t *pend_z = nullptr;
	/* per-view capacities: they start from what a view typically needs and follow the counters of the batches
	 * rendered so far (a batch that overflows one is re-run with a larger one), never the worst case of the map */
	int cap_cands = 0, cap_edges = 0, cap_blk_events = 0;
	bool caps_dirty = false;
	int vw_grid = 0;                         /* k_viswalk: resident blocks, persistent over views */
	/* the per-view constants reach the device through mapped pinned staging read by a kernel (two buffers, alternating), not
	 * through a copy engine: a host->device copy was seen to wait for the frames of the batch before */
	rc_view_t *h_views[2] = { nullptr, nullptr }; size_t cap_hviews = 0; int hviews_turn = 0;
	unsigned long long *h_counters_dev = nullptr; /* the mirror as the device sees it (mapped pinned memory) */
	unsigned long long *h_counters = nullptr; /* pinned mirror of the batch counters, copied at the end of every batch */
	rc_view_t *d_views = nullptr;
	void *d_work[24] = {};
	bool have_batch = false;
	void *d_flush = nullptr;

	/* entities of the next batches */
	uint16_t *d_alphamap = nullptr; uint8_t *d_addmap = nullptr, *d_modmap = nullptr;
	rc_item_t *d_items = nullptr; rc_corner_t *d_corners = nullptr; rc_sortent_t *d_sortents = nullptr;
	rc_glow_t *d_glow = nullptr; int nglow = 0, glow_views = 0;
	rc_ddecal_t *d_decals = nullptr; int32_t *d_decal_first = nullptr; uint32_t *d_decal_gen = nullptr; bool have_decals = false;
	int nitems = 0, ncorners = 0, nsortents = 0;
	std::vector<rc_item_t> h_items;
	int ntris = 0;                       /* triangles of the items */
	int nfast_items = 0, fast_tris = 0;  /* leading opaque run */
	unsigned long long *d_trikeys = nullptr; size_t cap_trikeys = 0;
	int32_t *d_biglist = nullptr; size_t cap_big = 0;
	rc_cmd2d_t *d_cmds2d = nullptr;
	int32_t *d_pairs = nullptr; size_t cap_pairs = 0; int fast_maxcount = 0;
	rc_bent_t *d_bents = nullptr; int32_t *d_bfaces = nullptr; int nbents = 0, nbfaces = 0;
	int32_t *d_leafkey = nullptr, *d_tbord = nullptr, *d_tbcounts = nullptr; rc_surf_t *d_tbsurfs = nullptr; rc_edge_t *d_tbedges = nullptr; int ntb = 0; size_t cap_tb = 0;
	int32_t *d_leafbase = nullptr, *d_viewbm = nullptr; uint32_t *d_bmvis = nullptr; size_t cap_leafbase = 0, cap_bmvis = 0, cap_viewbm = 0;
	int32_t *d_partbox = nullptr; size_t cap_partbox = 0; uint8_t *d_itemvis = nullptr; size_t cap_itemvis = 0;
	int32_t *d_item_order = nullptr; rc_poly_t *d_polys = nullptr;
	size_t cap_item_order = 0, cap_polys = 0;

	/* studio models */
	uint8_t *d_sstatic = nullptr; rc_spart_t *d_parts = nullptr; float *d_bones = nullptr, *d_blight = nullptr, *d_mvpos = nullptr; int32_t *d_mlight = nullptr;
	int nparts = 0, totalverts = 0, totalnorms = 0; size_t cap_mvpos = 0, cap_mlight = 0; bool studio_dirty = false;

	/* dynamic lights of the next batch */
	rc_dlight_t *d_dlights = nullptr;
	uint32_t *d_dlightbits = nullptr;
	int ndlights = 0;

	/* display tail */
	uint32_t *d_screenmap = nullptr;
	int screen_bpp = 0;
	void *d_display = nullptr;
	size_t display_bytes = 0;
} D;

bool ck( cudaError_t e, const char *what )
{
	if( e == cudaSuccess )
		return true;
	snprintf( D.err, sizeof( D.err ), "%s: %s", what, cudaGetErrorString( e ));
	return false;
}
#define CK( call ) do { if( !ck(( call ), #call )) return -1; } while( 0 )
#define KCHECK( name, st ) do { if( D.sync_each ) { cudaError_t e_ = cudaStreamSynchronize( st ); if( e_ == cudaSuccess ) e_ = cudaGetLastError(); if( e_ != cudaSuccess ) { snprintf( D.err, sizeof( D.err ), "kernel %s: %s", name, cudaGetErrorString( e_ )); return -1; } } } while( 0 )

template <typename T> int upload( T **dst, const T *src, size_t n )
{
	if( *dst ) { cudaFree( *dst ); *dst = nullptr; }
	size_t bytes = sizeof( T ) * ( n ? n : 1 );
	CK( cudaMalloc( (void **)dst, bytes ));
	if( n && src ) CK( cudaMemcpy( *dst, src, sizeof( T ) * n, cudaMemcpyHostToDevice ));
	else CK( cudaMemset( *dst, 0, bytes ));
	return 0;
}

/* per-batch state (entity lists, bones): the buffer is kept and only grows - cudaFree / cudaMalloc synchronise the whole
 * device, which would wait for the frames an earlier asynchronous batch is still copying to the host */
template <typename T> int upload_keep( T **dst, const T *src, size_t n )
{
	static std::map<void **, size_t> caps; /* capacity (elements) of the buffer behind every pointer slot */
	size_t &cap = caps[(void **)dst];
	if( !*dst ) cap = 0;
	if( n > cap || !*dst )
	{
		const size_t want = n + n / 2 + 16;
		if( *dst ) { cudaFree( *dst ); *dst = nullptr; }
		CK( cudaMalloc( (void **)dst, sizeof( T ) * want ));
		cap = want;
	}
	if( n && src ) CK( cudaMemcpy( *dst, src, sizeof( T ) * n, cudaMemcpyHostToDevice ));
	return 0;
}

template <typename T> int dalloc( T **dst, size_t n )
{
	if( *dst ) { cudaFree( *dst ); *dst = nullptr; }
	CK( cudaMalloc( (void **)dst, sizeof( T ) * ( n ? n : 1 )));
	return 0;
}

void free_batch()
{
	for( void *&p : D.d_work ) { if( p ) cudaFree( p ); p = nullptr; }
	if( D.d_views ) { cudaFree( D.d_views ); D.d_views = nullptr; }
	D.cap_views = 0; D.have_batch = false;
}

int ensure_batch( int nviews, int width, int height )
{
	if( nviews <= D.cap_views && width == D.cap_w && height == D.cap_h && !D.caps_dirty )
		return 0;
	const int keep = nviews > D.cap_views ? nviews : D.cap_views;
	const bool same_size = width == D.cap_w && height == D.cap_h;
	free_batch();
	const rc_scene_t &sc = D.sc;
	rc_work_t &w = D.w;
	int cap = same_size ? keep : nviews;
	w.width = width; w.height = height;
	const int worst_cands = sc.numsurfaces < 65000 ? ( sc.numsurfaces > 0 ? sc.numsurfaces : 1 ) : 65000;
	long long me = (long long)sc.numsurfedges + sc.numsurfaces + 64;
	const int worst_edges = (int32_t)( me < 2000000 ? me : 2000000 );
	if( !D.cap_cands ) D.cap_cands = worst_cands < 2048 ? worst_cands : 2048;
	if( D.cap_cands > worst_cands ) D.cap_cands = worst_cands;
	if( !D.cap_edges ) D.cap_edges = 6 * D.cap_cands + 64;
	if( D.cap_edges > worst_edges ) D.cap_edges = worst_edges;
	if( !D.cap_blk_events || !same_size ) D.cap_blk_events = SP_LINES * ( width <= 640 ? 192 : 384 );
	D.caps_dirty = false;
	w.max_cands = D.cap_cands;
	w.max_edges = D.cap_edges;
	w.visbytes_padded = ( sc.visbytes + 15 ) & ~15;
	w.nodewords = ( sc.numnodes + 31 ) / 32 + 1;
	w.leafwords = ( sc.numleafs + 31 ) / 32 + 1;
	w.surfwords = ( sc.numsurfaces + 31 ) / 32 + 1;
	w.stack_ints = (int32_t)rc_vis_scratch_ints2( sc.numnodes, sc.numleafs );
	{
		const size_t budget = 200 * 1024;
		size_t need = ((((size_t)( w.visbytes_padded >> 2 ) + w.nodewords + w.leafwords + w.surfwords ) * 4 ) + 15 ) & ~(size_t)15;
		w.vis_stage = 0;
		if( need + rc_vis_tree_bytes( &sc ) <= budget )
		{
			w.vis_stage |= VW_STAGE_TREE; need += rc_vis_tree_bytes( &sc );
			if( need + (size_t)w.stack_ints * 4 <= budget )
			{
				w.vis_stage |= VW_STAGE_SCRATCH; need += (size_t)w.stack_ints * 4;
				const size_t lists = (( 2 * ( (size_t)sc.numsurfaces + sc.nummarks )) + 15 ) & ~(size_t)15;
				if( need + lists <= budget )
				{
					w.vis_stage |= VW_STAGE_LISTS; need += lists;
				}
			}
		}
		if( const char *e = getenv( "RC_VIS_STAGE" )) /* measurement only: which of tree / scratch / lists k_viswalk keeps in shared memory */
		{
			const int want = atoi( e );
			size_t n2 = ((((size_t)( w.visbytes_padded >> 2 ) + w.nodewords + w.leafwords + w.surfwords ) * 4 ) + 15 ) & ~(size_t)15;
			int got = 0;
			if(( want & VW_STAGE_TREE ) && ( w.vis_stage & VW_STAGE_TREE )) { got |= VW_STAGE_TREE; n2 += rc_vis_tree_bytes( &sc ); }
			if(( want & VW_STAGE_SCRATCH ) && ( w.vis_stage & VW_STAGE_SCRATCH )) { got |= VW_STAGE_SCRATCH; n2 += (size_t)w.stack_ints * 4; }
			if(( want & VW_STAGE_LISTS ) && ( w.vis_stage & VW_STAGE_LISTS )) { got |= VW_STAGE_LISTS; n2 += (( 2 * ( (size_t)sc.numsurfaces + sc.nummarks )) + 15 ) & ~(size_t)15; }
			w.vis_stage = got; need = n2;
		}
		w.vis_smem = (int32_t)need;
	}
	{
		/* k_viswalk blocks are persistent over views: as many as the device keeps resident */
		int per_sm = 0;
		CK( cudaOccupancyMaxActiveBlocksPerMultiprocessor( &per_sm, k_viswalk<VW_NTHREADS>, VW_NTHREADS, (size_t)w.vis_smem ));
		if( per_sm < 1 ) per_sm = 1;
		/* a batch with many more views than resident blocks gains more from blocks than from a staged tree: keep only the
		 * walk's scratch in shared memory (measured, 4096 views of Map S: 0.96 -> 0.60 ms; 256 views: staged is faster,
		 * 49 vs 57 us - profiles/r03t_viswalk_staging.txt) */
		if( !getenv( "RC_VIS_STAGE" ) && ( w.vis_stage & VW_STAGE_SCRATCH ) && ( w.vis_stage & VW_STAGE_TREE ) && nviews >= 8 * per_sm * D.num_sms )
		{
			const size_t light = (((((size_t)( w.visbytes_padded >> 2 ) + w.nodewords + w.leafwords + w.surfwords ) * 4 ) + 15 ) & ~(size_t)15 ) + (size_t)w.stack_ints * 4;
			int per_sm2 = 0;
			CK( cudaOccupancyMaxActiveBlocksPerMultiprocessor( &per_sm2, k_viswalk<VW_NTHREADS>, VW_NTHREADS, light ));
			if( per_sm2 > per_sm )
			{
				w.vis_stage = VW_STAGE_SCRATCH; w.vis_smem = (int32_t)light; per_sm = per_sm2;
			}
		}
		D.vw_grid = per_sm * D.num_sms;
		D.vw_wide = getenv( "RC_VW_WIDE" ) ? atoi( getenv( "RC_VW_WIDE" )) != 0 : ( sc.numnodes >= 4096 );
	}

	int k = 0;
#define WALLOC( field, type, count ) do { type *p = nullptr; if( dalloc( &p, (size_t)( count )) != 0 ) return -1; w.field = p; D.d_work[k++] = p; } while( 0 )
	WALLOC( walkstack, int32_t, (size_t)D.vw_grid * RC_MAX_SLICES * w.stack_ints );
	WALLOC( cands, rc_cand_t, (size_t)cap * w.max_cands );
	WALLOC( candof, uint16_t, (size_t)cap * ( sc.numsurfaces + 2 ));
	WALLOC( facerecs, rc_facerec_t, (size_t)cap * w.max_cands );
	WALLOC( facewrote, uint8_t, (size_t)cap * w.max_cands );
	w.fix_cap = cap * 16 + 1024;
	WALLOC( fix_list, int32_t, (size_t)3 * w.fix_cap * RC_MAX_SLICES );
	WALLOC( surfs, rc_surf_t, (size_t)cap * w.max_cands );
	WALLOC( edges, rc_edge_t, (size_t)cap * w.max_edges );
	WALLOC( ncands, int32_t, cap );
	WALLOC( nedges, int32_t, cap );
	w.nbands = ( height + SP_LINES - 1 ) / SP_LINES;
	w.ev_ecap = width * height <= 640 * 480 ? EV_ECAP_SMALL : EV_ECAP_BIG;
	w.blk_events = D.cap_blk_events;
	w.view_events = (size_t)w.nbands * w.blk_events;
	w.scanbits = nullptr; w.scanbits_words = ( w.max_cands + 32 ) / 32;
	if( w.scanbits_words > SP_BITWORDS )
		WALLOC( scanbits, uint32_t, (size_t)RC_MAX_SLICES * LN_GRID * LN_WARPS * w.scanbits_words );
	WALLOC( bandsurf, uint16_t, (size_t)cap * w.nbands * ( (size_t)w.max_cands + 2 ));
	WALLOC( bandnloc, int32_t, (size_t)cap * w.nbands );
	WALLOC( events, unsigned long long, (size_t)cap * w.view_events );
	WALLOC( lineinfo, rc_lineinfo_t, (size_t)cap * height );
	WALLOC( heavy_list, int32_t, (size_t)cap * height );
	WALLOC( wide_list, int32_t, (size_t)cap * height );
	WALLOC( color, uint16_t, (size_t)cap * width * height );
	WALLOC( zbuf, int16_t, (size_t)cap * width * height );
#undef WALLOC
	if( dalloc( &D.d_views, cap ) != 0 ) return -1;
	CK( cudaMemset( w.color, 0, (size_t)cap * width * height * 2 ));
	CK( cudaMemset( w.zbuf, 0, (size_t)cap * width * height * 2 ));
	w.views = D.d_views;
	D.cap_views = cap; D.cap_w = width; D.cap_h = height;
	D.main_color = w.color; D.main_zbuf = w.zbuf; D.use_alt = false;
	if( D.alt_color ) { cudaFree( D.alt_color ); D.alt_color = nullptr; }
	if( D.alt_zbuf ) { cudaFree( D.alt_zbuf ); D.alt_zbuf = nullptr; }
	D.alt_px = 0;
	return 0;
}

/* after a batch: did a per-view capacity overflow?  Grow it from the counters (what the batch wanted, + 25 %) and tell
 * the caller to render again.  Returns 1 = re-run, 0 = fine, -1 = a limit that cannot grow was hit */
int check_capacities()
{
	const unsigned long long *c = D.h_counters;
	const rc_scene_t &sc = D.sc;
	int again = 0;
	if( c[RC_CTR_OVERFLOW_STACK] || c[RC_CTR_OVERFLOW_CACHE] )
	{
		snprintf( D.err, sizeof( D.err ), "batch dropped work: %llu views past the BSP depth limit, %llu surface cache blocks past the luxel limit",
			c[RC_CTR_OVERFLOW_STACK], c[RC_CTR_OVERFLOW_CACHE] );
		return -1;
	}
	if( c[RC_CTR_OVERFLOW_CANDS] )
	{
		const int worst = sc.numsurfaces < 65000 ? sc.numsurfaces : 65000;
		if( D.cap_cands >= worst ) { snprintf( D.err, sizeof( D.err ), "more than %d visible faces in a view", worst ); return -1; }
		long long want = (long long)c[RC_CTR_MAX_CANDS] * 5 / 4 + 64;
		D.cap_cands = (int)( want < worst ? want : worst );
		if( D.cap_edges < 6 * D.cap_cands + 64 ) D.cap_edges = 6 * D.cap_cands + 64;
		again = 1;
	}
	if( c[RC_CTR_OVERFLOW_EDGES] )
	{
		long long want = (long long)c[RC_CTR_MAX_EDGES] * 5 / 4 + 64;
		if( want <= D.cap_edges ) want = (long long)D.cap_edges * 2;
		if( want > 4000000 ) { snprintf( D.err, sizeof( D.err ), "more than 4M edges in a view" ); return -1; }
		D.cap_edges = (int)want;
		again = 1;
	}
	if( c[RC_CTR_OVERFLOW_EVENTS] )
	{
		long long want = (long long)c[RC_CTR_MAX_BAND_EVENTS] * 5 / 4 + 64;
		if( want <= D.cap_blk_events ) want = (long long)D.cap_blk_events * 2;
		if( want > ( 1 << 24 )) { snprintf( D.err, sizeof( D.err ), "more than 16M edge events in a band" ); return -1; }
		D.cap_blk_events = (int)(( want + 31 ) & ~31ll );
		again = 1;
	}
	if( again ) D.caps_dirty = true;
	return again;
}

/* the device stages of one batch on D.stream; ev != NULL -> stage boundaries are recorded */
/* the batch work record restricted to views [v0, v1): every per-view array starts at v0; slice k owns its own cache
 * work list, cursors and bit-set scratch so that slices can run on different streams at the same time */
rc_work_t make_slice( int k, int v0, int v1 )
{
	rc_work_t ws = D.w;
	const rc_scene_t &sc = D.sc;
	const size_t nslots = (size_t)sc.numsurfaces * 4;
	ws.nviews = v1 - v0;
	ws.views += v0;
	ws.walkstack += (size_t)k * D.vw_grid * ws.stack_ints; /* per resident block of the slice's k_viswalk */
	ws.cands += (size_t)v0 * ws.max_cands; ws.surfs += (size_t)v0 * ws.max_cands;
	ws.candof += (size_t)v0 * ( sc.numsurfaces + 2 );
	ws.facerecs += (size_t)v0 * ws.max_cands;
	ws.facewrote += (size_t)v0 * ws.max_cands;
	ws.fix_list += (size_t)3 * ws.fix_cap * k; ws.fix_count = D.d_nlist + 16 * k + 7;
	ws.edges += (size_t)v0 * ws.max_edges;
	ws.ncands += v0; ws.nedges += v0;
	ws.bandnloc += (size_t)v0 * ws.nbands;
	ws.bandsurf += (size_t)v0 * ws.nbands * ( (size_t)ws.max_cands + 2 );
	ws.events += (size_t)v0 * ws.view_events;
	ws.lineinfo += (size_t)v0 * ws.height;
	ws.color += (size_t)v0 * ws.width * ws.height; ws.zbuf += (size_t)v0 * ws.width * ws.height;
	if( ws.scanbits ) ws.scanbits += (size_t)k * LN_GRID * LN_WARPS * ws.scanbits_words;
	ws.cache_need = D.d_need + (size_t)k * ( nslots / 32 + 2 );
	ws.cache_keys = D.d_keys;
	ws.cache_list = D.d_list + (size_t)k * nslots;
	ws.cache_flist = D.d_flist + (size_t)k * nslots; ws.cache_units = D.d_units + (size_t)k * nslots * 32;
	ws.cache_lux = D.d_cachelux + (size_t)k * nslots * CACHE_FLUX; ws.cache_src = D.d_cachesrc + (size_t)k * nslots * CACHE_FLUX;
	ws.cache_fn = D.d_nlist + 16 * k + 10;
	ws.cache_nlist = D.d_nlist + 16 * k; ws.cache_cursor = ws.cache_nlist + 1; ws.line_cursor = ws.cache_nlist + 2; ws.heavy_count = ws.cache_nlist + 5; ws.wide_count = ws.cache_nlist + 6;
	ws.heavy_list += (size_t)v0 * ws.height;
	ws.wide_list += (size_t)v0 * ws.height;
	ws.styles = D.d_styles; ws.counters = D.d_counters;
	ws.dlights = D.d_dlights; ws.dlightbits = D.d_dlightbits; ws.ndlights = D.ndlights;
	ws.items = D.d_items; ws.corners = D.d_corners; ws.sortents = D.d_sortents; ws.nitems = D.nitems; ws.nsortents = D.nsortents;
	ws.nglow = D.nglow; ws.glow = D.nglow ? D.d_glow + (size_t)v0 * D.nglow : nullptr;
	ws.decals = D.have_decals ? D.d_decals : nullptr; ws.decal_first = D.have_decals ? D.d_decal_first : nullptr; ws.decal_gen = D.have_decals ? D.d_decal_gen : nullptr;
	ws.item_order = D.d_item_order ? D.d_item_order + (size_t)v0 * D.nitems : nullptr;
	ws.alphamap = D.d_alphamap; ws.addmap = D.d_addmap; ws.modmap = D.d_modmap;
	ws.polys = D.d_polys ? D.d_polys + (size_t)v0 * ( D.ntris - D.fast_tris ) : nullptr;
	ws.nfast_items = D.nfast_items; ws.fast_tris = D.fast_tris;
	ws.trikeys = D.d_trikeys ? D.d_trikeys + (size_t)v0 * ws.width * ws.height : nullptr;
	ws.big_cap = (int32_t)( D.cap_big / RC_MAX_SLICES ); ws.big_list = D.d_biglist ? D.d_biglist + (size_t)2 * ws.big_cap * k : nullptr;
	ws.big_count = D.d_nlist + 16 * k + 8; ws.pair_count = D.d_nlist + 16 * k + 9;
	ws.pairs = D.d_pairs ? D.d_pairs + (size_t)v0 * ( D.nfast_items > 0 ? D.nfast_items : 1 ) : nullptr;
	ws.bents = D.d_bents; ws.nbents = D.nbents; ws.bfaces = D.d_bfaces; ws.nbfaces = D.nbfaces;
	ws.leafbase = D.d_leafbase ? D.d_leafbase + (size_t)v0 * sc.numleafs : nullptr;
	ws.leafkey = D.d_leafkey ? D.d_leafkey + (size_t)v0 * sc.numleafs : nullptr;
	ws.tbord = D.d_tbord; ws.ntb = D.ntb;
	ws.tb_surfs = D.d_tbsurfs ? D.d_tbsurfs + (size_t)v0 * D.ntb * RC_TB_MAXSURFS : nullptr;
	ws.tb_edges = D.d_tbedges ? D.d_tbedges + (size_t)v0 * D.ntb * RC_TB_MAXEDGES : nullptr;
	ws.tb_counts = D.d_tbcounts ? D.d_tbcounts + (size_t)v0 * D.ntb * 2 : nullptr;
	ws.viewbm = D.d_viewbm ? D.d_viewbm + v0 : nullptr;
	ws.bmvis = D.d_bmvis ? D.d_bmvis + (size_t)v0 * ( ws.nodewords + ws.leafwords ) : nullptr;
	ws.partbox = D.d_partbox; ws.itemvis = D.d_itemvis ? D.d_itemvis + (size_t)v0 * D.nitems : nullptr;
	ws.sstatic = D.d_sstatic; ws.parts = D.d_parts; ws.bones = D.d_bones; ws.blight = D.d_blight; ws.mvpos = D.d_mvpos; ws.mlight = D.d_mlight;
	return ws;
}

/* the studio models' vertices and light levels do not depend on the view: once per batch */
int launch_studio_points( cudaStream_t st )
{
	if( !D.nparts || !D.nitems || !D.studio_dirty )
		return 0;
	const rc_work_t ws = make_slice( 0, 0, D.w.nviews );
	const int np = D.totalverts + D.totalnorms;
	k_partbox_init<<<( D.nparts * 6 + 255 ) / 256, 256, 0, st>>>( ws, D.nparts ); D.launches++;
	k_studio_points<<<( np + 255 ) / 256, 256, 0, st>>>( D.sc, ws, D.nparts, D.totalverts, D.totalnorms ); D.launches++;
	KCHECK( "k_studio_points", st );
	D.studio_dirty = false;
	return 0;
}

/* all stages of one slice on stream st; evs != NULL: stage boundaries are recorded there */
int launch_slice( const rc_work_t &ws, int k, cudaStream_t st, cudaEvent_t *evs, bool draw )
{
	const rc_scene_t &sc = D.sc;
	fill32( ws.cache_need, 0u, (size_t)(( sc.numsurfaces * 4 + 31 ) / 32 + 1 ), st );
	fill32( ws.cache_nlist, 0u, 16, st );
	if( D.nbents ) fill32( ws.viewbm, 0u, (size_t)ws.nviews, st );
	if( D.nbents ) fill32( ws.leafbase, 0xFFFFFFFFu, (size_t)ws.nviews * sc.numleafs, st );
	if( evs ) CK( cudaEventRecord( evs[0], st ));
	/* a handful of views of a large map (C3: 8 views, 15 k nodes, levels thousands of nodes wide): 1024 threads per view */
	if( D.vw_wide && ws.nviews <= D.num_sms )
		k_viswalk<1024><<<ws.nviews < D.vw_grid ? ws.nviews : D.vw_grid, 1024, (size_t)ws.vis_smem, st>>>( sc, ws );
	else
	k_viswalk<VW_NTHREADS><<<ws.nviews < D.vw_grid ? ws.nviews : D.vw_grid, VW_NTHREADS, (size_t)ws.vis_smem, st>>>( sc, ws );
	D.launches++;
	KCHECK( "k_viswalk", st );
	if( evs ) CK( cudaEventRecord( evs[1], st ));
	{
		/* x = view: the blocks that have candidates (the first chunks of every view) are issued first, ahead of the
		 * thousands of empty ones (max_cands is the face count of the map) */
		dim3 g( ws.nviews, ( ws.max_cands + FC_THREADS - 1 ) / FC_THREADS );
		k_faces<<<g, FC_THREADS, 0, st>>>( sc, ws ); D.launches++;
		KCHECK( "k_faces", st );
		k_faces_fix<<<D.num_sms * 2, FC_THREADS, 0, st>>>( sc, ws ); D.launches++;
		KCHECK( "k_faces_fix", st );
		if( D.nbents && D.nbfaces )
		{
			dim3 gb(( D.nbfaces + 63 ) / 64, ws.nviews );
			k_bmodels<<<gb, 64, 0, st>>>( sc, ws ); D.launches++;
			KCHECK( "k_bmodels", st );
		}
		if( D.ntb )
		{
			k_tbrush_pieces<<<( ws.nviews * D.nbents + 63 ) / 64, 64, 0, st>>>( sc, ws ); D.launches++;
			KCHECK( "k_tbrush_pieces", st );
		}
	}
	if( evs ) CK( cudaEventRecord( evs[2], st ));
	{
		/* the surface cache is only read by k_draw: its kernels run on a side stream, beside span generation */
		cudaStream_t aux = D.cache_aside ? D.aux_stream[k & 1] : st; /* as many side streams as slice streams: every stream costs a hardware queue */
		CK( cudaEventRecord( D.ev_cache[k][0], st ));
		if( D.cache_aside ) CK( cudaStreamWaitEvent( aux, D.ev_cache[k][0], 0 ));
		int nslots = sc.numsurfaces * 4;
		k_cache_list<<<( nslots + 255 ) / 256, 256, 0, aux>>>( sc, ws ); D.launches++;
		k_cache_light<<<148 * 4, 256, 0, aux>>>( sc, ws ); D.launches++;
		k_cache_texels<<<148 * 8, 256, 0, aux>>>( sc, ws ); D.launches++;
		k_cache_publish<<<( nslots + 255 ) / 256, 256, 0, aux>>>( sc, ws ); D.launches++;
		k_cache_build<<<148 * 2, CACHE_THREADS, 0, aux>>>( sc, ws ); D.launches++; /* ragged blocks only: few or none */
		KCHECK( "k_cache_list/build", aux );
		CK( cudaEventRecord( D.ev_cache[k][1], aux ));
	}
	if( evs ) CK( cudaEventRecord( evs[3], st ));
	const int nlines = ws.nviews * ws.height;
	{
		dim3 g( ws.nviews, ws.nbands ); /* x = view: blocks are issued band by band, see k_events */
		/* wide frames of large maps: more threads per band block (few blocks, thousands of edges each) */
		const int evt = D.ev_threads ? D.ev_threads : ( ws.width > 640 ? 1024 : 256 );
		if( evt == 1024 ) k_events<1024><<<g, 1024, rc_events_smem( ws.ev_ecap, ws.scanbits_words ), st>>>( ws );
		else if( evt == 512 ) k_events<512><<<g, 512, rc_events_smem( ws.ev_ecap, ws.scanbits_words ), st>>>( ws );
		else k_events<256><<<g, 256, rc_events_smem( ws.ev_ecap, ws.scanbits_words ), st>>>( ws );
		D.launches++;
		KCHECK( "k_events", st );
		const int lblocks = ( nlines + LN_WARPS - 1 ) / LN_WARPS;
		k_lines<<<lblocks < LN_GRID ? lblocks : LN_GRID, LN_WARPS * 32, 0, st>>>( ws ); D.launches++;
		KCHECK( "k_lines", st );
		/* the few lines with more events / band surfaces than k_lines takes (none on small maps: the blocks find an empty list) */
		k_lines_wide<<<WD_GRID, WD_THREADS, 0, st>>>( ws ); D.launches++;
		KCHECK( "k_lines_wide", st );
	}
	if( evs ) CK( cudaEventRecord( evs[4], st ));
	if( D.cache_aside ) CK( cudaStreamWaitEvent( st, D.ev_cache[k][1], 0 ));
	if( draw )
	{
		const int lpw = rc_draw_lpw( ws.width, ws.height );
		const int dblocks = ( nlines / lpw + DR_WARPS - 1 ) / DR_WARPS;
		k_draw<<<dblocks < DR_GRID ? dblocks : DR_GRID, DR_WARPS * 32, rc_draw_smem( ws.width, lpw ), st>>>( sc, ws, 0, nlines, lpw );
		D.launches++;
		KCHECK( "k_draw", st );
		if( evs ) CK( cudaEventRecord( evs[5], st ));
		if( D.nitems && ( D.ntris || D.ntb ))
		{
			/* entities: after the world's spans are in the frame (R_DrawEntitiesOnList follows R_EdgeDrawing, r_main.c:1090) */
			if( launch_studio_points( st ) != 0 ) return -1;
			k_item_cull<<<( ws.nviews * D.nitems + 127 ) / 128, 128, 0, st>>>( ws ); D.launches++;
			if( D.fast_tris )
			{
				k_tri_keys_init<<<D.num_sms * 8, 256, 0, st>>>( ws ); D.launches++;
				/* x = slot in the pair list (the live ones first; the rest of the upper bound exits at once), y = 128 triangles */
				dim3 gf( ws.nviews * D.nfast_items, ( D.fast_maxcount + TF_THREADS * TF_ROUNDS - 1 ) / ( TF_THREADS * TF_ROUNDS ));
				k_tris_fast<<<gf, TF_THREADS, 0, st>>>( ws ); D.launches++;
				KCHECK( "k_tris_fast", st );
				k_tris_fast_big<<<D.num_sms * 8, 128, 0, st>>>( ws ); D.launches++;
				k_tri_resolve<<<D.num_sms * 16, 128, 0, st>>>( sc, ws ); D.launches++;
				KCHECK( "k_tri_resolve", st );
			}
			const int nord = D.ntris - D.fast_tris;
			if( nord > 0 || D.ntb )
			{
				k_item_order<<<ws.nviews, 128, 0, st>>>( ws ); D.launches++;
				if( nord > 0 )
				{
					dim3 gp(( nord + 127 ) / 128, ws.nviews );
					k_tri_polys<<<gp, 128, 0, st>>>( ws, nord ); D.launches++;
				}
				dim3 gt( ws.nviews, D.tr_strips );
				k_tris_ordered<<<gt, TR_THREADS, 0, st>>>( sc, ws, nord ); D.launches++;
			}
			KCHECK( "k_tris", st );
		}
	}
	else if( evs ) CK( cudaEventRecord( evs[5], st ));
	if( evs ) CK( cudaEventRecord( evs[6], st ));
	return 0;
}

/* the stages of slices [k0, k0 + kn) of a batch, forked from st onto the slice streams and joined back into it */
int enqueue_segment( const rc_work_t *wss, int k0, int kn, cudaStream_t st )
{
	CK( cudaEventRecord( D.ev_fork, st ));
	for( int j = 0; j < kn; j++ )
	{
		cudaStream_t ss = D.slice_stream[j];
		CK( cudaStreamWaitEvent( ss, D.ev_fork, 0 ));
		if( launch_slice( wss[j], k0 + j, ss, nullptr, true ) != 0 ) return -1;
		CK( cudaEventRecord( D.ev_join[j], ss ));
	}
	for( int j = 0; j < kn; j++ ) CK( cudaStreamWaitEvent( st, D.ev_join[j], 0 ));
	return 0;
}

/* everything the launches of a segment are made from */
unsigned long long segment_key( const rc_work_t *wss, int kn )
{
	unsigned long long h = 1469598103934665603ull;
	auto mix = [&h]( const void *p, size_t n ) { const unsigned char *b = (const unsigned char *)p; for( size_t i = 0; i < n; i++ ) { h ^= b[i]; h *= 1099511628211ull; } };
	mix( &D.sc, sizeof( D.sc ));
	for( int j = 0; j < kn; j++ ) mix( &wss[j], sizeof( rc_work_t ));
	const long long scal[] = { D.vw_grid, D.cache_aside, D.nbents, D.nbfaces, D.ntb, D.nitems, D.ntris, D.fast_tris, D.nfast_items, D.fast_maxcount, D.num_sms, kn };
	mix( scal, sizeof( scal ));
	return h ? h : 1;
}

/*
 * The device stages of the batch.  nslices == 1: one stream, stage boundaries optionally recorded in D.ev (the
 * per-kernel timings of the bench).  nslices > 1: the views are cut into slices, slice k on stream k % nstreams;
 * with host outputs (color_out / z_out) every slice's frames are copied out on D.copy as soon as the slice is drawn.
 */
int run_batch( int nslices, int nstreams, bool timed, uint16_t *color_out, int16_t *z_out )
{
	rc_work_t &w = D.w;
	cudaStream_t st = D.stream;
	const size_t vpx = (size_t)w.width * w.height;
	fill32( D.d_counters, 0u, 2 * RC_CTR_COUNT, st );
	if( nslices > 1 && launch_studio_points( st ) != 0 ) return -1; /* ahead of the fork: every slice reads the result */
	if( nslices <= 1 )
	{
		const rc_work_t ws = make_slice( 0, 0, w.nviews );
		if( launch_slice( ws, 0, st, timed ? D.ev : nullptr, true ) != 0 ) return -1;
		if( color_out ) CK( cudaMemcpyAsync( color_out, w.color, (size_t)w.nviews * vpx * 2, cudaMemcpyDeviceToHost, st ));
		if( z_out ) CK( cudaMemcpyAsync( z_out, w.zbuf, (size_t)w.nviews * vpx * 2, cudaMemcpyDeviceToHost, st ));
		k_counters_out<<<1, RC_CTR_COUNT, 0, st>>>( D.d_counters, D.h_counters_dev );
		CK( cudaGetLastError());
		return 0;
	}
	/*
	 * The slices in pairs ("segments"): the two slices of a segment run side by side on the two slice streams, and the
	 * frames of a finished segment leave on the copy stream while the next one is computed.  The ~30 launches of a segment
	 * are replayed from a CUDA graph: while a copy engine is busy every single launch of the process takes longer (measured:
	 * the stages of a batch 1.8x slower beside another stream's device->host copy; a 60-kernel graph 1.07x; a kernel that is
	 * already running is not affected), and a batch with host outputs always has copies in flight.  A graph is keyed by a
	 * hash of every launch argument (scene and slice records, the counts the grids are made from): any change re-captures it.
	 */
	if( timed ) CK( cudaEventRecord( D.ev[0], st ));
	const bool graphs = D.use_graphs && !timed && !D.sync_each && nstreams == 2 && !( nslices & 1 );
	for( int k0 = 0; k0 < nslices; k0 += 2 )
	{
		const int kn = k0 + 2 <= nslices ? 2 : 1;
		int v0s[2], v1s[2];
		rc_work_t wss[2];
		for( int j = 0; j < kn; j++ )
		{
			v0s[j] = (int)((long long)w.nviews * ( k0 + j ) / nslices ); v1s[j] = (int)((long long)w.nviews * ( k0 + j + 1 ) / nslices );
			wss[j] = make_slice( k0 + j, v0s[j], v1s[j] );
		}
		bool replayed = false;
		if( graphs )
		{
			Dev::batch_graph_t &g = D.graphs[( k0 >> 1 ) & 3][D.use_alt ? 1 : 0];
			const unsigned long long key = segment_key( wss, kn );
			if( g.exec && g.key == key )
				D.launches += g.launches;
			else
			{
				if( g.exec ) { cudaGraphExecDestroy( g.exec ); g.exec = nullptr; }
				const unsigned long long l0 = D.launches;
				cudaGraph_t graph = nullptr;
				bool ok = cudaStreamBeginCapture( st, cudaStreamCaptureModeThreadLocal ) == cudaSuccess;
				if( ok )
				{
					ok = enqueue_segment( wss, k0, kn, st ) == 0;
					ok = ( cudaStreamEndCapture( st, &graph ) == cudaSuccess ) && ok && graph;
				}
				if( ok ) ok = cudaGraphInstantiate( &g.exec, graph, 0 ) == cudaSuccess;
				if( graph ) cudaGraphDestroy( graph );
				if( !ok )
				{
					/* never seen; the plain path still works */
					cudaGetLastError();
					g.exec = nullptr; D.use_graphs = false; D.launches = l0;
				}
				else { g.key = key; g.launches = D.launches - l0; }
			}
			if( g.exec )
			{
				CK( cudaGraphLaunch( g.exec, st ));
				replayed = true;
			}
		}
		if( !replayed && enqueue_segment( wss, k0, kn, st ) != 0 ) return -1;
		if( color_out || z_out )
		{
			CK( cudaEventRecord( D.ev_slice[k0], st ));
			CK( cudaStreamWaitEvent( D.copy, D.ev_slice[k0], 0 ));
			const size_t a0 = (size_t)v0s[0] * vpx, n = ( (size_t)v1s[kn - 1] - v0s[0] ) * vpx;
			if( color_out ) CK( cudaMemcpyAsync( color_out + a0, w.color + a0, n * 2, cudaMemcpyDeviceToHost, D.copy ));
			if( z_out ) CK( cudaMemcpyAsync( z_out + a0, w.zbuf + a0, n * 2, cudaMemcpyDeviceToHost, D.copy ));
		}
	}
	if( timed ) CK( cudaEventRecord( D.ev[RC_BENCH_STAGES], st ));
	k_counters_out<<<1, RC_CTR_COUNT, 0, st>>>( D.d_counters, D.h_counters_dev );
	CK( cudaGetLastError());
	return 0;
}

} /* namespace */

extern "C" {

int rc_dev_init( int device )
{
	/* a batch with host outputs keeps 6 streams busy (stages of two slices, their surface-cache side streams, the copy stream,
	 * the batch stream); the default of 8 hardware queues per device is shared with whatever else lives in the process, and
	 * two streams on one queue serialise (measured: the second pair of slices ran 1.4x slower behind the first pair's
	 * copies).  Only effective when the CUDA context does not exist yet; a host process that creates it first sets this itself */
	setenv( "CUDA_DEVICE_MAX_CONNECTIONS", "32", 0 );
	int n = 0;
	if( !ck( cudaGetDeviceCount( &n ), "cudaGetDeviceCount" ) || n <= 0 )
	{
		if( !D.err[0] ) snprintf( D.err, sizeof( D.err ), "no CUDA device" );
		return -1;
	}
	if( device < 0 || device >= n ) device = 0;
	CK( cudaSetDevice( device ));
	cudaDeviceProp prop;
	CK( cudaGetDeviceProperties( &prop, device ));
	snprintf( D.name, sizeof( D.name ), "%s (sm_%d%d, %d SMs)", prop.name, prop.major, prop.minor, prop.multiProcessorCount );
	D.device = device; D.num_sms = prop.multiProcessorCount;
	D.sync_each = getenv( "RC_SYNC_EACH" ) != nullptr;
	CK( cudaHostAlloc( (void **)&D.h_counters, sizeof( unsigned long long ) * RC_CTR_COUNT, cudaHostAllocMapped ));
	CK( cudaHostGetDevicePointer( (void **)&D.h_counters_dev, D.h_counters, 0 ));
	memset( D.h_counters, 0, sizeof( unsigned long long ) * RC_CTR_COUNT );
	CK( cudaStreamCreateWithFlags( &D.stream, cudaStreamNonBlocking ));
	CK( cudaStreamCreateWithFlags( &D.copy, cudaStreamNonBlocking ));
	for( auto &e : D.ev_slice ) CK( cudaEventCreateWithFlags( &e, cudaEventDisableTiming ));
	CK( cudaEventCreateWithFlags( &D.ev_fork, cudaEventDisableTiming ));
	for( auto &e : D.ev_join ) CK( cudaEventCreateWithFlags( &e, cudaEventDisableTiming ));
	D.use_graphs = getenv( "RC_NO_GRAPHS" ) == nullptr;
	if( const char *e = getenv( "RC_EV_THREADS" )) { const int n = atoi( e ); if( n == 256 || n == 512 || n == 1024 ) D.ev_threads = n; }
	if( const char *e = getenv( "RC_TR_STRIPS" )) { const int n = atoi( e ); if( n >= 1 && n <= 64 ) D.tr_strips = n; }
	for( auto &ss : D.slice_stream ) CK( cudaStreamCreateWithFlags( &ss, cudaStreamNonBlocking ));
	for( auto &ss : D.aux_stream ) CK( cudaStreamCreateWithFlags( &ss, cudaStreamNonBlocking ));
	for( auto &e : D.ev_cache ) { CK( cudaEventCreate( &e[0] )); CK( cudaEventCreate( &e[1] )); }
	for( auto &e : D.ev ) CK( cudaEventCreate( &e ));
	CK( cudaFuncSetAttribute( k_viswalk<VW_NTHREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024 + 64 ));
	CK( cudaFuncSetAttribute( k_viswalk<1024>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024 + 64 ));
	CK( cudaFuncSetAttribute( k_events<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rc_events_smem( EV_ECAP_BIG, 65536 / 32 + 2 )));
	CK( cudaFuncSetAttribute( k_events<512>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rc_events_smem( EV_ECAP_BIG, 65536 / 32 + 2 )));
	CK( cudaFuncSetAttribute( k_events<1024>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rc_events_smem( EV_ECAP_BIG, 65536 / 32 + 2 )));
	CK( cudaFuncSetAttribute( k_draw, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rc_draw_smem( 8192, 1 )));
	for( int i = 0; i < 256; i++ ) D.h_styles[i] = 256;
	if( upload( &D.d_styles, D.h_styles, 256 ) != 0 ) return -1;
	if( dalloc( &D.d_counters, RC_CTR_COUNT ) != 0 ) return -1;
	CK( cudaMemset( D.d_counters, 0, sizeof( unsigned long long ) * RC_CTR_COUNT ));
	if( dalloc( &D.d_nlist, 16 * RC_MAX_SLICES ) != 0 ) return -1;
	D.ready = true;
	return 0;
}

void rc_dev_shutdown( void )
{
	if( !D.ready ) return;
	cudaDeviceSynchronize();
	free_batch();
	/* everything rc_dev_init / the uploads created, so that R_Shutdown + R_Init (vid_restart) starts clean */
	void **ptrs[] = { (void **)&D.d_leafkey, (void **)&D.d_tbord, (void **)&D.d_tbcounts, (void **)&D.d_tbsurfs, (void **)&D.d_tbedges, (void **)&D.d_bents, (void **)&D.d_bfaces, (void **)&D.d_leafbase, (void **)&D.d_viewbm, (void **)&D.d_bmvis, (void **)&D.d_pairs, (void **)&D.d_partbox, (void **)&D.d_itemvis, (void **)&D.d_cmds2d, (void **)&D.d_trikeys, (void **)&D.d_biglist, (void **)&D.d_sstatic, (void **)&D.d_parts, (void **)&D.d_bones, (void **)&D.d_blight, (void **)&D.d_mvpos, (void **)&D.d_mlight, (void **)&D.d_item_order, (void **)&D.d_polys, (void **)&D.d_alphamap, (void **)&D.d_addmap, (void **)&D.d_modmap, (void **)&D.d_items, (void **)&D.d_glow, (void **)&D.d_decals, (void **)&D.d_decal_first, (void **)&D.d_decal_gen, (void **)&D.d_corners, (void **)&D.d_sortents, (void **)&D.d_display, (void **)&D.d_screenmap, (void **)&D.d_dlights, (void **)&D.d_dlightbits,
		(void **)&D.d_colormap, (void **)&D.d_lightgamma, (void **)&D.d_sintable, (void **)&D.d_textures, (void **)&D.d_texels,
		(void **)&D.d_styles, (void **)&D.d_keys, (void **)&D.d_need, (void **)&D.d_list, (void **)&D.d_nlist, (void **)&D.d_flist, (void **)&D.d_units, (void **)&D.d_cachelux, (void **)&D.d_cachesrc,
		(void **)&D.d_counters, &D.d_flush };
	for( void **p : ptrs ) { if( *p ) cudaFree( *p ); *p = nullptr; }
	for( void *&p : D.d_scene ) { if( p ) cudaFree( p ); p = nullptr; }
	if( D.h_counters ) { cudaFreeHost( D.h_counters ); D.h_counters = nullptr; }
	for( auto &h : D.h_views ) { if( h ) cudaFreeHost( h ); h = nullptr; }
	D.cap_hviews = 0;
	D.cap_cands = D.cap_edges = D.cap_blk_events = 0; D.caps_dirty = false;
	D.nitems = D.ncorners = D.nsortents = 0; D.h_items.clear(); D.ntris = 0; D.cap_item_order = D.cap_polys = 0;
	D.nparts = D.totalverts = D.totalnorms = 0; D.cap_mvpos = D.cap_mlight = 0; D.studio_dirty = false;
	D.nfast_items = D.fast_tris = 0; D.cap_trikeys = D.cap_big = 0; D.cap_partbox = D.cap_itemvis = D.cap_pairs = 0; D.nbents = D.nbfaces = 0; D.cap_leafbase = D.cap_bmvis = D.cap_viewbm = 0; D.ntb = 0; D.cap_tb = 0;
	D.display_bytes = 0; D.ndlights = 0; D.screen_bpp = 0; D.cache_texels = 0; D.ntextures = 0;
	D.sc = rc_scene_t{}; D.h_surfaces.clear(); D.h_textures.clear();
	cudaStream_t *streams[] = { &D.stream, &D.copy };
	for( cudaStream_t *st : streams ) { if( *st ) cudaStreamDestroy( *st ); *st = nullptr; }
	for( auto &ss : D.slice_stream ) { if( ss ) cudaStreamDestroy( ss ); ss = nullptr; }
	for( auto &ss : D.aux_stream ) { if( ss ) cudaStreamDestroy( ss ); ss = nullptr; }
	for( auto &e : D.ev_slice ) { if( e ) cudaEventDestroy( e ); e = nullptr; }
	for( auto &e : D.ev ) { if( e ) cudaEventDestroy( e ); e = nullptr; }
	for( auto &e : D.ev_cache ) for( auto &x : e ) { if( x ) cudaEventDestroy( x ); x = nullptr; }
	if( D.ev_fork ) { cudaEventDestroy( D.ev_fork ); D.ev_fork = nullptr; }
	for( auto &e : D.ev_join ) { if( e ) cudaEventDestroy( e ); e = nullptr; }
	for( auto &row : D.graphs ) for( auto &g : row ) { if( g.exec ) cudaGraphExecDestroy( g.exec ); g.exec = nullptr; g.key = 0; }
	D.launches = 0;
	D.ready = false;
}

const char *rc_dev_last_error( void ) { return D.err; }
const char *rc_dev_name( void ) { return D.name; }
uint64_t rc_dev_launch_count( void ) { return D.launches; }
uint64_t rc_dev_h2d_bytes( int reset ) { const uint64_t v = D.h2d_bytes; if( reset ) D.h2d_bytes = 0; return v; }

int rc_dev_upload_luts( const uint16_t *colormap, const uint16_t *lightgamma, const int32_t *sintable )
{
	if( upload( &D.d_colormap, colormap, 32 * 8192 ) != 0 ) return -1;
	if( upload( &D.d_lightgamma, lightgamma, 1024 ) != 0 ) return -1;
	if( upload( &D.d_sintable, sintable, 1280 ) != 0 ) return -1;
	D.sc.colormap = D.d_colormap; D.sc.lightgamma = D.d_lightgamma; D.sc.sintable = D.d_sintable;
	return rc_dev_flush_cache();
}

/* A19 on the device: block row = (texture, plane), threads stride over the plane's texels */
__global__ void __launch_bounds__( 256 ) k_tex_encode( const rc_texture_t *tex, const rc_texsrc_t *src, const uint32_t *rgba, uint16_t *texels )
{
	const int ti = blockIdx.y / 5, plane = blockIdx.y - ti * 5;
	const rc_texture_t t = tex[ti];
	if( t.ofs[0] < 0 )
		return;
	const int level = plane == 4 ? 0 : plane;
	const int32_t ofs = plane == 4 ? t.alpha_ofs : t.ofs[plane];
	if( ofs < 0 )
		return;
	const int32_t n = max( 1, t.width >> level ) * max( 1, t.height >> level );
	const rc_texsrc_t sr = src[ti];
	const uint32_t *image = rgba + sr.rgba_ofs;
	for( int32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x )
		texels[(size_t)ofs + k] = rc_tex_encode( image, t.width, t.height, sr.flags, plane, k );
}

int rc_dev_upload_textures( const rc_texture_t *tex, const rc_texsrc_t *src, int ntex, const uint32_t *rgba, size_t nrgba, size_t ntexels )
{
	CK( cudaStreamSynchronize( D.stream ));
	if( upload( &D.d_textures, tex, (size_t)ntex ) != 0 ) return -1;
	if( dalloc( &D.d_texels, ntexels ) != 0 ) return -1;
	CK( cudaMemsetAsync( D.d_texels, 0, sizeof( uint16_t ) * ntexels, D.stream ));
	D.h_textures.assign( tex, tex + ntex );
	D.ntextures = ntex;
	D.sc.textures = D.d_textures; D.sc.texels = D.d_texels; D.sc.numtextures = ntex;
	if( ntex > 0 && nrgba > 0 )
	{
		/* the RGBA sources live on the device only while they are encoded */
		rc_texsrc_t *d_src = nullptr; uint32_t *d_rgba = nullptr;
		if( upload( &d_src, src, (size_t)ntex ) != 0 || upload( &d_rgba, rgba, nrgba ) != 0 ) { cudaFree( d_src ); cudaFree( d_rgba ); return -1; }
		int32_t maxn = 1;
		for( int i = 0; i < ntex; i++ ) if( tex[i].ofs[0] >= 0 && tex[i].width * tex[i].height > maxn ) maxn = tex[i].width * tex[i].height;
		int bx = ( maxn + 255 ) / 256; if( bx > 64 ) bx = 64;
		for( int t0 = 0; t0 < ntex; t0 += 8192 ) /* gridDim.y <= 65535 */
		{
			const int nt = ntex - t0 < 8192 ? ntex - t0 : 8192;
			k_tex_encode<<<dim3( bx, nt * 5 ), 256, 0, D.stream>>>( D.d_textures + t0, d_src + t0, d_rgba, D.d_texels ); D.launches++;
		}
		cudaError_t e = cudaStreamSynchronize( D.stream );
		cudaFree( d_src ); cudaFree( d_rgba );
		if( e != cudaSuccess ) { snprintf( D.err, sizeof( D.err ), "k_tex_encode: %s", cudaGetErrorString( e )); return -1; }
	}
	return rc_dev_flush_cache();
}

int rc_dev_upload_scene( const rc_scene_t *sc, size_t cache_texels )
{
	CK( cudaStreamSynchronize( D.stream ));
	free_batch();
	D.cap_cands = D.cap_edges = D.cap_blk_events = 0; D.caps_dirty = false;
	rc_scene_t d = *sc;
	rc_plane_t *planes = (rc_plane_t *)D.d_scene[0]; rc_node_t *nodes = (rc_node_t *)D.d_scene[1];
	rc_leaf_t *leafs = (rc_leaf_t *)D.d_scene[2]; rc_surface_t *surfaces = (rc_surface_t *)D.d_scene[3];
	int32_t *marks = (int32_t *)D.d_scene[4], *surfedges = (int32_t *)D.d_scene[5];
	uint16_t *edges = (uint16_t *)D.d_scene[6]; float *verts = (float *)D.d_scene[7];
	uint8_t *visdata = (uint8_t *)D.d_scene[8], *lightdata = (uint8_t *)D.d_scene[9];
	uint16_t *cache = (uint16_t *)D.d_scene[10];
	if( upload( &planes, sc->planes, (size_t)sc->numplanes ) != 0 ) return -1;
	if( upload( &nodes, sc->nodes, (size_t)sc->numnodes ) != 0 ) return -1;
	if( upload( &leafs, sc->leafs, (size_t)sc->numleafs ) != 0 ) return -1;
	if( upload( &surfaces, sc->surfaces, (size_t)sc->numsurfaces ) != 0 ) return -1;
	if( upload( &marks, sc->marksurfaces, (size_t)sc->nummarks ) != 0 ) return -1;
	if( upload( &surfedges, sc->surfedges, (size_t)sc->numsurfedges ) != 0 ) return -1;
	if( upload( &edges, sc->edges, (size_t)sc->numedges * 2 ) != 0 ) return -1;
	if( upload( &verts, sc->verts, (size_t)sc->numverts * 3 ) != 0 ) return -1;
	if( upload( &visdata, sc->visdata, (size_t)sc->visdata_len ) != 0 ) return -1;
	if( upload( &lightdata, sc->lightdata, (size_t)sc->lightdata_len ) != 0 ) return -1;
	if( upload( &cache, (const uint16_t *)nullptr, cache_texels ) != 0 ) return -1;

	/* derived tables: every leaf's PVS row expanded once (Mod_DecompressPVS, engine/common/mod_bmodel.c:1054: zero
	 * runs are {0, count} pairs), and the node each surface hangs off */
	uint32_t *pvsdense = (uint32_t *)D.d_scene[11]; int32_t *surfnode = (int32_t *)D.d_scene[12]; uint32_t *surfback = (uint32_t *)D.d_scene[13];
	{
		const int32_t vpad = ( sc->visbytes + 15 ) & ~15;
		std::vector<uint8_t> dense((size_t)( sc->numleafs > 0 ? sc->numleafs : 1 ) * vpad, 0 );
		for( int32_t l = 0; l < sc->numleafs; l++ )
		{
			uint8_t *out = dense.data() + (size_t)l * vpad;
			const int32_t ofs = sc->leafs[l].visofs;
			if( ofs < 0 || !sc->visdata_len )
			{
				memset( out, 0xFF, (size_t)sc->visbytes );
				continue;
			}
			const uint8_t *in = sc->visdata + ofs, *end = sc->visdata + sc->visdata_len;
			int32_t d = 0;
			while( d < sc->visbytes && in < end )
			{
				if( *in ) out[d++] = *in++;
				else { d += in[1]; in += 2; }
			}
		}
		if( upload( &pvsdense, (const uint32_t *)dense.data(), dense.size() / 4 ) != 0 ) return -1;
		std::vector<int32_t> sn((size_t)( sc->numsurfaces > 0 ? sc->numsurfaces : 1 ), 0 );
		for( int32_t n = 0; n < sc->numnodes; n++ )
			for( int32_t k = 0; k < sc->nodes[n].numsurfaces; k++ )
				if( sc->nodes[n].firstsurface + k < sc->numsurfaces )
					sn[sc->nodes[n].firstsurface + k] = n;
		if( upload( &surfnode, sn.data(), sn.size()) != 0 ) return -1;
		std::vector<uint32_t> sb((size_t)( sc->numsurfaces + 31 ) / 32 + 2, 0u );
		for( int32_t i = 0; i < sc->numsurfaces; i++ )
			if( sc->surfaces[i].flags & RC_SURF_PLANEBACK ) sb[i >> 5] |= 1u << ( i & 31 );
		if( upload( &surfback, sb.data(), sb.size()) != 0 ) return -1;
	}
	float *sedgeverts = (float *)D.d_scene[14];
	int32_t *sedgepartner = (int32_t *)D.d_scene[15];
	{
		std::vector<float> sv((size_t)( sc->numsurfedges > 0 ? sc->numsurfedges : 1 ) * 6, 0.0f );
		std::vector<int32_t> sp((size_t)( sc->numsurfedges > 0 ? sc->numsurfedges : 1 ), -1 );
		rc_build_sedge_tables( sc, sv.data(), sp.data());
		if( upload( &sedgeverts, sv.data(), sv.size()) != 0 ) return -1;
		if( upload( &sedgepartner, sp.data(), sp.size()) != 0 ) return -1;
	}
	D.d_scene[15] = sedgepartner; d.sedgepartner = sedgepartner;
	D.d_scene[11] = pvsdense; D.d_scene[12] = surfnode; D.d_scene[13] = surfback; D.d_scene[14] = sedgeverts;
	d.pvsdense = pvsdense; d.surfnode = surfnode; d.surfback = surfback; d.sedgeverts = sedgeverts;
	D.d_scene[0] = planes; D.d_scene[1] = nodes; D.d_scene[2] = leafs; D.d_scene[3] = surfaces; D.d_scene[4] = marks;
	D.d_scene[5] = surfedges; D.d_scene[6] = edges; D.d_scene[7] = verts; D.d_scene[8] = visdata; D.d_scene[9] = lightdata;
	D.d_scene[10] = cache;
	d.planes = planes; d.nodes = nodes; d.leafs = leafs; d.surfaces = surfaces; d.marksurfaces = marks;
	d.surfedges = surfedges; d.edges = edges; d.verts = verts; d.visdata = visdata; d.lightdata = lightdata;
	d.cache = cache;
	d.textures = D.d_textures; d.texels = D.d_texels; d.numtextures = D.ntextures;
	d.colormap = D.d_colormap; d.lightgamma = D.d_lightgamma; d.sintable = D.d_sintable;
	D.sc = d;
	D.cache_texels = cache_texels;
	D.h_surfaces.assign( sc->surfaces, sc->surfaces + sc->numsurfaces );
	D.have_decals = false; /* the tables are per face of the previous world */

	size_t nslots = (size_t)sc->numsurfaces * 4;
	if( upload( &D.d_keys, (const rc_cachekey_t *)nullptr, nslots ) != 0 ) return -1;
	if( upload( &D.d_dlightbits, (const uint32_t *)nullptr, (size_t)sc->numsurfaces + 1 ) != 0 ) return -1;
	D.ndlights = 0;
	if( upload( &D.d_need, (const uint32_t *)nullptr, RC_MAX_SLICES * ( nslots / 32 + 2 )) != 0 ) return -1;
	if( upload( &D.d_list, (const int32_t *)nullptr, RC_MAX_SLICES * nslots ) != 0 ) return -1;
	if( dalloc( &D.d_flist, (size_t)RC_MAX_SLICES * nslots ) != 0 || dalloc( &D.d_units, (size_t)RC_MAX_SLICES * nslots * 32 ) != 0 ) return -1;
	if( dalloc( &D.d_cachelux, (size_t)RC_MAX_SLICES * nslots * CACHE_FLUX ) != 0 || dalloc( &D.d_cachesrc, (size_t)RC_MAX_SLICES * nslots * CACHE_FLUX ) != 0 ) return -1;
	return 0;
}

int rc_dev_upload_blendluts( const uint16_t *alphamap, const uint8_t *addmap, const uint8_t *modmap )
{
	if( upload( &D.d_alphamap, alphamap, (size_t)3 * 1024 * 256 ) != 0 ) return -1;
	if( upload( &D.d_addmap, addmap, 65536 ) != 0 ) return -1;
	if( upload( &D.d_modmap, modmap, 65536 ) != 0 ) return -1;
	return 0;
}

int rc_dev_set_items( const rc_item_t *items, int nitems, const rc_corner_t *corners, int ncorners, const rc_sortent_t *sortents, int nsortents )
{
	if( !nitems && !D.nitems ) return 0;
	CK( cudaStreamSynchronize( D.stream ));
	D.nitems = nitems; D.nsortents = nsortents; D.ncorners = ncorners;
	if( !nitems ) { D.ntris = 0; return 0; }
	if( upload_keep( &D.d_items, items, (size_t)nitems ) != 0 ) return -1;
	if( upload_keep( &D.d_corners, corners, (size_t)ncorners ) != 0 ) return -1;
	if( upload_keep( &D.d_sortents, sortents, (size_t)nsortents ) != 0 ) return -1;
	D.h_items.assign( items, items + nitems );
	D.ntris = 0;
	for( int i = 0; i < nitems; i++ ) D.ntris += items[i].count;
	/* the leading run of opaque, fixed-order items takes the order-free path when it is worth its key buffer */
	D.nfast_items = D.fast_tris = 0;
	for( int i = 0; i < nitems && items[i].drawer == RC_DRAW_FILL && items[i].sortent < 0; i++ ) { D.nfast_items++; D.fast_tris += items[i].count; }
	if( D.fast_tris < 64 || D.nfast_items > 65535 ) D.nfast_items = D.fast_tris = 0;
	D.fast_maxcount = 0;
	for( int i = 0; i < D.nfast_items; i++ ) if( items[i].count > D.fast_maxcount ) D.fast_maxcount = items[i].count;
	D.h2d_bytes += sizeof( rc_item_t ) * (size_t)nitems + sizeof( rc_corner_t ) * (size_t)ncorners + sizeof( rc_sortent_t ) * (size_t)nsortents;
	return 0;
}

int rc_dev_set_decals( const rc_ddecal_t *recs, int nrecs, const int32_t *face_first_count, const uint32_t *face_gen, int nfaces )
{
	if( nfaces != D.sc.numsurfaces ) { snprintf( D.err, sizeof( D.err ), "decal tables for %d faces, the world has %d", nfaces, D.sc.numsurfaces ); return -1; }
	if( rc_dev_wait() != 0 ) return -1;
	CK( cudaDeviceSynchronize());
	if( upload_keep( &D.d_decals, recs, (size_t)( nrecs > 0 ? nrecs : 1 )) != 0 ) return -1;
	if( upload_keep( &D.d_decal_first, face_first_count, (size_t)nfaces * 2 ) != 0 ) return -1;
	if( upload_keep( &D.d_decal_gen, face_gen, (size_t)nfaces ) != 0 ) return -1;
	D.have_decals = true;
	D.h2d_bytes += sizeof( rc_ddecal_t ) * (size_t)nrecs + 12 * (size_t)nfaces;
	return 0;
}

/* the per-view records of the batch's kRenderGlow sprites, [view][sprite]; the next render call must bring `nviews` views */
int rc_dev_set_glow( const rc_glow_t *recs, int nglow, int nviews )
{
	D.nglow = 0; D.glow_views = 0;
	if( nglow <= 0 || nviews <= 0 ) return 0;
	CK( cudaStreamSynchronize( D.stream ));
	if( upload_keep( &D.d_glow, recs, (size_t)nglow * nviews ) != 0 ) return -1;
	D.nglow = nglow; D.glow_views = nviews;
	D.h2d_bytes += sizeof( rc_glow_t ) * (size_t)nglow * nviews;
	return 0;
}

int rc_dev_set_bents( const rc_bent_t *bents, int n )
{
	if( n > RC_MAX_BENTS ) n = RC_MAX_BENTS;
	D.nbents = 0; D.nbfaces = 0; D.ntb = 0;
	if( n <= 0 ) return 0;
	std::vector<int32_t> bf;
	for( int e = 0; e < n; e++ )
		for( int f = 0; !bents[e].trans && f < bents[e].numfaces && f < 65536; f++ ) bf.push_back(( e << 16 ) | f );
	if( upload_keep( &D.d_bents, bents, (size_t)n ) != 0 ) return -1;
	if( upload_keep( &D.d_bfaces, bf.data(), bf.size()) != 0 ) return -1;
	D.h2d_bytes += sizeof( rc_bent_t ) * (size_t)n + 4 * bf.size();
	std::vector<int32_t> ord( n );
	int ntb = 0;
	for( int e = 0; e < n; e++ ) ord[e] = bents[e].trans ? ntb++ : -1;
	if( upload_keep( &D.d_tbord, ord.data(), ord.size()) != 0 ) return -1;
	D.ntb = ntb;
	D.nbents = n; D.nbfaces = (int)bf.size();
	return 0;
}

int rc_dev_draw2d( const rc_cmd2d_t *cmds, int n, int view )
{
	if( !D.have_batch || view < 0 || view >= D.w.nviews ) { snprintf( D.err, sizeof( D.err ), "rc_dev_draw2d: no such frame" ); return -1; }
	if( n <= 0 ) return 0;
	if( upload( &D.d_cmds2d, cmds, (size_t)n ) != 0 ) return -1;
	const rc_work_t ws = make_slice( 0, 0, D.w.nviews );
	for( int i = 0; i < n; i++ )
	{
		const int px = cmds[i].w * cmds[i].height;
		if( px <= 0 ) continue;
		int blocks = ( px + 255 ) / 256;
		if( blocks > D.num_sms * 8 ) blocks = D.num_sms * 8;
		k_2d<<<blocks, 256, 0, D.stream>>>( D.sc, ws, D.d_cmds2d, i, view ); D.launches++;
	}
	KCHECK( "k_2d", D.stream );
	CK( cudaStreamSynchronize( D.stream ));
	return 0;
}

int rc_dev_upload_studio_static( const void *arena, size_t bytes )
{
	CK( cudaStreamSynchronize( D.stream ));
	if( upload( &D.d_sstatic, (const uint8_t *)arena, bytes + 16 > 16 ? bytes : 16 ) != 0 ) return -1;
	D.h2d_bytes += bytes;
	return 0;
}

int rc_dev_set_studio( const rc_spart_t *parts, int nparts, const float *bones, int nbones, const float *blightvecs, int totalverts, int totalnorms )
{
	if( !nparts && !D.nparts ) return 0;
	CK( cudaStreamSynchronize( D.stream ));
	D.nparts = nparts; D.totalverts = totalverts; D.totalnorms = totalnorms;
	if( !nparts ) return 0;
	if( upload_keep( &D.d_parts, parts, (size_t)nparts ) != 0 ) return -1;
	if( upload_keep( &D.d_bones, bones, (size_t)nbones * 12 ) != 0 ) return -1;
	if( upload_keep( &D.d_blight, blightvecs, (size_t)nbones * 3 ) != 0 ) return -1;
	if( (size_t)totalverts * 4 > D.cap_mvpos ) { if( dalloc( &D.d_mvpos, (size_t)totalverts * 4 ) != 0 ) return -1; D.cap_mvpos = (size_t)totalverts * 4; } /* float4 per vertex */
	if( (size_t)totalnorms > D.cap_mlight ) { if( dalloc( &D.d_mlight, (size_t)totalnorms ) != 0 ) return -1; D.cap_mlight = (size_t)totalnorms; }
	D.h2d_bytes += sizeof( rc_spart_t ) * (size_t)nparts + sizeof( float ) * 15 * (size_t)nbones;
	if( (size_t)nparts * 6 > D.cap_partbox ) { if( dalloc( &D.d_partbox, (size_t)nparts * 6 + 6 ) != 0 ) return -1; D.cap_partbox = (size_t)nparts * 6; }
	D.studio_dirty = true;
	return 0;
}

int rc_dev_update_surface_texnums( const int32_t *texnum, int n )
{
	CK( cudaStreamSynchronize( D.stream ));
	if( n > (int)D.h_surfaces.size()) n = (int)D.h_surfaces.size();
	for( int i = 0; i < n; i++ ) D.h_surfaces[i].texnum = texnum[i];
	CK( cudaMemcpy( (void *)D.sc.surfaces, D.h_surfaces.data(), sizeof( rc_surface_t ) * D.h_surfaces.size(), cudaMemcpyHostToDevice ));
	return 0;
}

int rc_dev_set_lightstyles( const int32_t *values )
{
	if( !memcmp( D.h_styles, values, sizeof( D.h_styles )))
		return 0;
	memcpy( D.h_styles, values, sizeof( D.h_styles ));
	CK( cudaStreamSynchronize( D.stream ));
	CK( cudaMemcpy( D.d_styles, D.h_styles, sizeof( D.h_styles ), cudaMemcpyHostToDevice ));
	D.h2d_bytes += sizeof( D.h_styles );
	return 0;
}

int rc_dev_set_dlights( const rc_dlight_t *lights, int n )
{
	if( n < 0 || n > RC_MAX_DLIGHTS ) return -1;
	if( !n )
	{
		D.ndlights = 0;
		return 0;
	}
	if( !D.sc.numsurfaces ) { snprintf( D.err, sizeof( D.err ), "rc_dev_set_dlights: no world" ); return -1; }
	CK( cudaStreamSynchronize( D.stream ));
	if( !D.d_dlights && dalloc( &D.d_dlights, RC_MAX_DLIGHTS ) != 0 ) return -1;
	CK( cudaMemcpyAsync( D.d_dlights, lights, sizeof( rc_dlight_t ) * n, cudaMemcpyHostToDevice, D.stream ));
	D.h2d_bytes += sizeof( rc_dlight_t ) * (size_t)n;
	k_dlight_mark<<<( D.sc.numsurfaces + 127 ) / 128, 128, 0, D.stream>>>( D.sc, D.d_dlights, n, D.d_dlightbits );
	D.launches++;
	CK( cudaGetLastError());
	D.ndlights = n;
	return 0;
}

int rc_dev_flush_cache( void )
{
	if( D.d_keys && D.sc.numsurfaces )
	{
		CK( cudaStreamSynchronize( D.stream ));
		CK( cudaMemset( D.d_keys, 0, sizeof( rc_cachekey_t ) * (size_t)D.sc.numsurfaces * 4 ));
	}
	return 0;
}

/* the per-batch buffers of the entity stages, sized for the batch capacity */
static int ensure_entity_buffers( int width, int height )
{
	if( D.nbents )
	{
		const size_t need_l = (size_t)D.cap_views * D.sc.numleafs, need_b = (size_t)D.cap_views * ( D.w.nodewords + D.w.leafwords );
		if( need_l > D.cap_leafbase ) { if( dalloc( &D.d_leafbase, need_l ) != 0 || dalloc( &D.d_leafkey, need_l ) != 0 ) return -1; D.cap_leafbase = need_l; }
		const size_t need_t = (size_t)D.cap_views * D.ntb;
		if( need_t > D.cap_tb )
		{
			if( dalloc( &D.d_tbsurfs, need_t * RC_TB_MAXSURFS ) != 0 || dalloc( &D.d_tbedges, need_t * RC_TB_MAXEDGES ) != 0 || dalloc( &D.d_tbcounts, need_t * 2 ) != 0 ) return -1;
			D.cap_tb = need_t;
		}
		if( need_b > D.cap_bmvis ) { if( dalloc( &D.d_bmvis, need_b ) != 0 ) return -1; D.cap_bmvis = need_b; }
		if( (size_t)D.cap_views > D.cap_viewbm ) { if( dalloc( &D.d_viewbm, (size_t)D.cap_views ) != 0 ) return -1; D.cap_viewbm = D.cap_views; }
	}
	if( D.nitems && ( D.ntris || D.ntb ))
	{
		const size_t need_v = (size_t)D.cap_views * D.nitems;
		if( need_v > D.cap_itemvis ) { if( dalloc( &D.d_itemvis, need_v ) != 0 ) return -1; D.cap_itemvis = need_v; }
		const size_t need_o = (size_t)D.cap_views * D.nitems, need_p = (size_t)D.cap_views * ( D.ntris - D.fast_tris ) + 1;
		if( D.fast_tris )
		{
			const size_t need_k = (size_t)D.cap_views * width * height, need_b = (size_t)RC_MAX_SLICES * ( (size_t)D.cap_views * 4096 + 65536 );
			if( need_k > D.cap_trikeys ) { if( dalloc( &D.d_trikeys, need_k ) != 0 ) return -1; D.cap_trikeys = need_k; }
			if( need_b > D.cap_big ) { if( dalloc( &D.d_biglist, 2 * need_b ) != 0 ) return -1; D.cap_big = need_b; }
			const size_t need_pr = (size_t)D.cap_views * D.nfast_items;
			if( need_pr > D.cap_pairs ) { if( dalloc( &D.d_pairs, need_pr ) != 0 ) return -1; D.cap_pairs = need_pr; }
		}
		if( need_o > D.cap_item_order ) { if( dalloc( &D.d_item_order, need_o ) != 0 ) return -1; D.cap_item_order = need_o; }
		if( need_p > D.cap_polys ) { if( dalloc( &D.d_polys, need_p ) != 0 ) return -1; D.cap_polys = need_p; }
	}
	return 0;
}

/* the batch's per-view constants to the device; buffers zeroed where the views leave a border */
static int begin_batch( const rc_view_t *views, int nviews, int width, int height, int clear )
{
	rc_work_t &w = D.w;
	if( D.nglow && D.glow_views != nviews ) { snprintf( D.err, sizeof( D.err ), "glow records cover %d views, the batch has %d", D.glow_views, nviews ); return -1; }
	w.nviews = nviews; w.clear = clear;
	cudaStream_t st = D.stream;
	if( (size_t)nviews > D.cap_hviews )
	{
		CK( cudaStreamSynchronize( st ));
		for( auto &h : D.h_views ) { if( h ) cudaFreeHost( h ); h = nullptr; }
		const size_t cap = (size_t)nviews > (size_t)D.cap_views ? (size_t)nviews : (size_t)D.cap_views;
		for( auto &h : D.h_views ) CK( cudaHostAlloc( (void **)&h, sizeof( rc_view_t ) * cap, cudaHostAllocMapped ));
		D.cap_hviews = cap;
	}
	{
		static_assert( sizeof( rc_view_t ) % 4 == 0, "rc_view_t is copied in 32-bit words" );
		rc_view_t *stage = D.h_views[D.hviews_turn ^= 1];
		memcpy( stage, views, sizeof( rc_view_t ) * nviews );
		uint32_t *stage_dev = nullptr;
		CK( cudaHostGetDevicePointer( (void **)&stage_dev, stage, 0 ));
		const size_t nwords = sizeof( rc_view_t ) / 4 * (size_t)nviews;
		k_copy32<<<(unsigned)(( nwords + 255 ) / 256 < 592 ? ( nwords + 255 ) / 256 : 592 ), 256, 0, st>>>( (uint32_t *)D.d_views, stage_dev, nwords );
	}
	D.h2d_bytes += sizeof( rc_view_t ) * (size_t)nviews;
	size_t px = (size_t)nviews * width * height;
	if( clear )
	{
		/* views that do not cover the whole buffer leave a border the span pass never touches */
		bool full = true;
		for( int i = 0; i < nviews && full; i++ )
			full = views[i].vrect[0] == 0 && views[i].vrect[1] == 0 && views[i].vrect[2] == width && views[i].vrect[3] == height;
		if( !full )
		{
			fill32( w.color, 0u, ( px + 1 ) / 2, st ); /* the buffers hold cap_views frames of an even pixel count or end on allocation slack */
			fill32( w.zbuf, 0u, ( px + 1 ) / 2, st );
		}
	}
	return 0;
}

int rc_dev_wait( void );

int rc_dev_render( const rc_view_t *views, int nviews, int width, int height, int clear, uint16_t *color_out, int16_t *z_out )
{
	if( !D.ready ) { snprintf( D.err, sizeof( D.err ), "device not initialised" ); return -1; }
	if( !D.sc.numsurfaces || !D.sc.surfaces ) { snprintf( D.err, sizeof( D.err ), "no world uploaded" ); return -1; }
	if( nviews <= 0 || width <= 0 || height <= 0 || height > 1200 ) { snprintf( D.err, sizeof( D.err ), "bad batch geometry" ); return -1; }
	if( rc_dev_wait() != 0 ) return -1; /* an asynchronous batch still in flight finishes first */
	for( int attempt = 0; ; attempt++ )
	{
		if( ensure_batch( nviews, width, height ) != 0 ) return -1;
		D.w.color = D.main_color; D.w.zbuf = D.main_zbuf; D.use_alt = false;
		if( ensure_entity_buffers( width, height ) != 0 ) return -1;
		if( begin_batch( views, nviews, width, height, clear ) != 0 ) return -1;
		cudaStream_t st = D.stream;
		{
			/* frames that go back to the host: 4 slices of views on two streams, so that the copies of the first pair
			 * overlap the stages of the second (the copies are the long pole: PCIe).  Frames that stay in HBM: one
			 * stream - running the slices side by side was measured and gains nothing (every kernel already exposes the
			 * whole batch's parallelism; the slices move in lockstep) */
			const bool out = color_out || z_out;
			const int nslices = out && nviews >= 16 * RC_MAX_SLICES ? RC_MAX_SLICES : 1;
			if( run_batch( nslices, 2, false, color_out, z_out ) != 0 ) return -1;
			if( out && nslices > 1 ) CK( cudaStreamSynchronize( D.copy ));
		}
		CK( cudaStreamSynchronize( st ));
		D.have_batch = true;
		/* a per-view capacity overflowed: nothing of this batch is trusted - grow it and render again */
		const int again = check_capacities();
		if( again < 0 ) return -1;
		if( !again ) return 0;
		if( attempt >= 6 ) { snprintf( D.err, sizeof( D.err ), "capacities still overflow after %d attempts", attempt + 1 ); return -1; }
	}
}


/*
 * Asynchronous batches.  rc_dev_render_async enqueues everything rc_dev_render does - stages per view slice, the frames of
 * every slice copied to the host on the copy stream - and returns; the frames are in color_out / z_out after rc_dev_wait.
 * Successive asynchronous batches draw into alternating sets of frame buffers, so the copies of batch k (PCIe, the long
 * pole) run under the stages of batch k + 1.  Before a new batch is enqueued the kernels (not the copies) of the previous
 * one are waited for and its overflow counters checked; a batch that overflowed a capacity is rendered again, synchronously.
 */
static int settle_pending()
{
	if( !D.pending )
		return 0;
	CK( cudaEventSynchronize( D.ev_kdone ));
	D.pending = false;
	const int again = check_capacities();
	if( again < 0 ) return -1;
	if( !again ) return 0;
	CK( cudaStreamSynchronize( D.copy ));
	std::vector<rc_view_t> v;
	v.swap( D.pend_views );
	return rc_dev_render( v.data(), (int)v.size(), D.pend_w, D.pend_h, D.pend_clear, D.pend_color, D.pend_z );
}

int rc_dev_wait( void )
{
	if( !D.ready ) return 0;
	if( settle_pending() != 0 ) return -1;
	CK( cudaStreamSynchronize( D.copy ));
	CK( cudaStreamSynchronize( D.stream ));
	return 0;
}

int rc_dev_render_async( const rc_view_t *views, int nviews, int width, int height, int clear, uint16_t *color_out, int16_t *z_out )
{
	if( !D.ready ) { snprintf( D.err, sizeof( D.err ), "device not initialised" ); return -1; }
	/* nothing to overlap (frames stay in HBM, or depend on the previous batch's pixels, or too few views to slice) */
	if( !clear || ( !color_out && !z_out ) || nviews < 16 * RC_MAX_SLICES || width <= 0 || height <= 0 || height > 1200 || !D.sc.numsurfaces )
		return rc_dev_render( views, nviews, width, height, clear, color_out, z_out );
	if( settle_pending() != 0 ) return -1;
	if( nviews > D.cap_views || width != D.cap_w || height != D.cap_h || D.caps_dirty )
	{
		if( rc_dev_wait() != 0 ) return -1; /* buffers are about to be replaced: nothing may be in flight */
		if( ensure_batch( nviews, width, height ) != 0 ) return -1;
	}
	if( ensure_entity_buffers( width, height ) != 0 ) return -1;
	const size_t px = (size_t)D.cap_views * width * height;
	if( D.alt_px < px )
	{
		if( rc_dev_wait() != 0 ) return -1;
		if( dalloc( &D.alt_color, px ) != 0 || dalloc( &D.alt_zbuf, px ) != 0 ) return -1;
		D.alt_px = px;
	}
	D.use_alt = !D.use_alt;
	D.w.color = D.use_alt ? D.alt_color : D.main_color;
	D.w.zbuf = D.use_alt ? D.alt_zbuf : D.main_zbuf;
	if( begin_batch( views, nviews, width, height, clear ) != 0 ) return -1;
	if( run_batch( RC_MAX_SLICES, 2, false, color_out, z_out ) != 0 ) return -1;
	if( !D.ev_kdone ) CK( cudaEventCreateWithFlags( &D.ev_kdone, cudaEventDisableTiming ));
	CK( cudaEventRecord( D.ev_kdone, D.stream ));
	D.pend_views.assign( views, views + nviews );
	D.pend_w = width; D.pend_h = height; D.pend_clear = clear; D.pend_color = color_out; D.pend_z = z_out;
	D.pending = true;
	D.have_batch = true;
	return 0;
}

int rc_dev_bench_resident( int iters, int flags, float *ms_total, float *ms_stage )
{
	if( rc_dev_wait() != 0 ) return -1;
	if( !D.have_batch ) { snprintf( D.err, sizeof( D.err ), "no batch rendered yet" ); return -1; }
	float tot = 0, stage[RC_BENCH_STAGES] = { 0 };
	bool overlapped = false;
	const size_t flush_bytes = (size_t)256 << 20; /* > 126 MB of L2 */
	if(( flags & RC_BENCH_FLUSH_L2 ) && !D.d_flush )
		CK( cudaMalloc( &D.d_flush, flush_bytes ));
	D.cache_aside = !( flags & RC_BENCH_SERIAL_CACHE );
	for( int it = 0; it < iters; it++ )
	{
		if( flags & RC_BENCH_FLUSH_L2 ) /* untimed: the events bracket the stages only */
			CK( cudaMemsetAsync( D.d_flush, it & 0xFF, flush_bytes, D.stream ));
		if( D.nparts ) D.studio_dirty = true; /* skinning is per-frame work (R_StudioDrawPoints): part of every iteration */
		if( flags & RC_BENCH_REBUILD_CACHE )
			CK( cudaMemsetAsync( D.d_keys, 0, sizeof( rc_cachekey_t ) * (size_t)D.sc.numsurfaces * 4, D.stream ));
		{
			const bool overlap = ( flags & RC_BENCH_OVERLAP ) && D.w.nviews >= 16 * RC_MAX_SLICES;
			if( run_batch( overlap ? RC_MAX_SLICES : 1, 2, true, nullptr, nullptr ) != 0 ) { D.cache_aside = true; return -1; }
			if( overlap ) overlapped = true;
		}
		CK( cudaStreamSynchronize( D.stream ));
		float ms;
		CK( cudaEventElapsedTime( &ms, D.ev[0], D.ev[RC_BENCH_STAGES] )); tot += ms;
		if( !overlapped )
		{
			for( int s = 0; s < RC_BENCH_STAGES; s++ ) { CK( cudaEventElapsedTime( &ms, D.ev[s], D.ev[s + 1] )); stage[s] += ms; }
			/* the cache kernels run beside span generation: their own events (the stream's events see nothing of them) */
			if( D.cache_aside ) { CK( cudaEventElapsedTime( &ms, D.ev_cache[0][0], D.ev_cache[0][1] )); stage[2] += ms; }
		}
	}
	D.cache_aside = true;
	if( ms_total ) *ms_total = tot / ( iters > 0 ? iters : 1 );
	if( ms_stage ) for( int s = 0; s < RC_BENCH_STAGES; s++ ) ms_stage[s] = stage[s] / ( iters > 0 ? iters : 1 );
	return 0;
}

int rc_dev_read_cache_block( int surf, int mip, uint16_t *out, int maxtexels, int *width, int *height, int *valid )
{
	if( surf < 0 || surf >= (int)D.h_surfaces.size() || mip < 0 || mip > 3 ) return -1;
	const rc_surface_t *sf = &D.h_surfaces[surf];
	int bw = sf->lightextents[0] >> mip, bh = sf->lightextents[1] >> mip;
	*width = bw; *height = bh; *valid = 0;
	if( sf->cache_ofs[mip] < 0 ) return 0;
	rc_cachekey_t k;
	CK( cudaMemcpy( &k, D.d_keys + surf * 4 + mip, sizeof( k ), cudaMemcpyDeviceToHost ));
	bool stale = !k.valid || k.texnum != sf->texnum; /* a block built with dynamic lights is a valid picture of the last batch */
	for( int j = 0; j < 4 && sf->styles[j] != 255 && !stale; j++ )
		if( k.light[j] != D.h_styles[sf->styles[j]] ) stale = true;
	*valid = !stale;
	if( *valid && bw * bh <= maxtexels )
		CK( cudaMemcpy( out, D.sc.cache + sf->cache_ofs[mip], (size_t)bw * bh * 2, cudaMemcpyDeviceToHost ));
	return 0;
}

int rc_dev_set_screenmap( const uint32_t *table, int bpp )
{
	if( bpp != 2 && bpp != 4 ) { snprintf( D.err, sizeof( D.err ), "rc_dev_set_screenmap: bpp %d not supported (2 or 4)", bpp ); return -1; }
	if( !D.d_screenmap && dalloc( &D.d_screenmap, 65536 ) != 0 ) return -1;
	CK( cudaMemcpy( D.d_screenmap, table, sizeof( uint32_t ) * 65536, cudaMemcpyHostToDevice ));
	D.screen_bpp = bpp;
	return 0;
}

int rc_dev_blit( int first, int nviews, void *dst, size_t stride, float *ms_kernel )
{
	if( !D.have_batch || !D.d_screenmap || first < 0 || nviews < 0 || first + nviews > D.w.nviews )
	{
		snprintf( D.err, sizeof( D.err ), "rc_dev_blit: no batch / no screen map / views out of range" );
		return -1;
	}
	const size_t fpx = (size_t)D.w.width * D.w.height, npx = fpx * nviews, bytes = npx * D.screen_bpp;
	if( !npx ) return 0;
	if( bytes > D.display_bytes )
	{
		if( D.d_display ) cudaFree( D.d_display );
		D.d_display = nullptr; D.display_bytes = 0;
		CK( cudaMalloc( &D.d_display, bytes ));
		D.display_bytes = bytes;
	}
	const uint16_t *src = D.w.color + fpx * first;
	const size_t want = ( npx / 8 + 255 ) / 256;
	const int grid = (int)( want < 148 * 8 ? ( want ? want : 1 ) : 148 * 8 );
	if( ms_kernel ) CK( cudaEventRecord( D.ev[8], D.stream ));
	if( D.screen_bpp == 2 ) k_blit<2><<<grid, 256, 0, D.stream>>>( src, D.d_screenmap, D.d_display, npx );
	else k_blit<4><<<grid, 256, 0, D.stream>>>( src, D.d_screenmap, D.d_display, npx );
	D.launches++;
	if( ms_kernel ) CK( cudaEventRecord( D.ev[9], D.stream ));
	CK( cudaGetLastError());
	if( dst )
	{
		const size_t rowbytes = (size_t)D.w.width * D.screen_bpp;
		if( stride < (size_t)D.w.width ) { snprintf( D.err, sizeof( D.err ), "rc_dev_blit: stride smaller than the frame width" ); return -1; }
		CK( cudaMemcpy2DAsync( dst, stride * D.screen_bpp, D.d_display, rowbytes, rowbytes, (size_t)D.w.height * nviews, cudaMemcpyDeviceToHost, D.stream ));
	}
	CK( cudaStreamSynchronize( D.stream ));
	if( ms_kernel ) CK( cudaEventElapsedTime( ms_kernel, D.ev[8], D.ev[9] ));
	return 0;
}

/* ---------------------------------------------------------------------------------------------- */
/* k_blit without an allocation (RefCuda_BlitPixelsNoAlloc, include/ref_cuda.h).
 *
 * rc_dev_blit keeps its screen map and its output staging buffer in cudaMalloc'ed memory.  Callers that must not allocate
 * (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc) get the same kernel over a file-scope device
 * array of a fixed size: the 256 KB screen map, the frame's pixel_t codes and the converted frame.  Every GPU that loads the
 * library holds the array (RC_STATIC_BLIT_BYTES, ~12.7 MB) until exit.  All calls share it and it keeps the last call's
 * contents; nothing in it has to start at zero (the table and the input are written before the kernel reads them, the
 * kernel writes every output pixel the copy-out reads).  Calls that overlap on the GPU (other threads, other streams) race.
 *
 * In host code g_static_blit is a host-side stand-in for the symbol, not a device address: the address the kernel gets
 * comes from cudaGetSymbolAddress, resolved once per device with that device current.                                   */

struct rc_static_blit_t
{
	uint32_t lut[65536];                    /* R_BuildScreenMap's table for the call's display format */
	uint16_t in[RC_STATIC_BLIT_MAXPX];      /* pixel_t codes */
	uint32_t out[RC_STATIC_BLIT_MAXPX];     /* display pixels (bpp 2: the first half is used) */
};
#define RC_STATIC_BLIT_BYTES sizeof( rc_static_blit_t )
static __device__ __align__( 256 ) unsigned char g_static_blit[RC_STATIC_BLIT_BYTES];
#define RC_STATIC_MAXDEV 64
static rc_static_blit_t *g_static_blit_ptr[RC_STATIC_MAXDEV]; /* device address of g_static_blit on every device that asked */

int rc_dev_blit_static( const uint32_t *table, int bpp, const uint16_t *pixels, size_t npx, void *dst )
{
	if( !table || !pixels || !dst || ( bpp != 2 && bpp != 4 )) { snprintf( D.err, sizeof( D.err ), "rc_dev_blit_static: bad arguments (bpp must be 2 or 4)" ); return -1; }
	if( npx == 0 ) return 0;
	if( npx > RC_STATIC_BLIT_MAXPX )
	{
		snprintf( D.err, sizeof( D.err ), "rc_dev_blit_static: %zu pixels, the static scratch holds %d (one 1920x1080 frame)", npx, RC_STATIC_BLIT_MAXPX );
		return -1;
	}
	int dev = -1;
	CK( cudaGetDevice( &dev ));
	if( dev < 0 || dev >= RC_STATIC_MAXDEV ) { snprintf( D.err, sizeof( D.err ), "rc_dev_blit_static: device ordinal %d out of range", dev ); return -1; }
	if( !g_static_blit_ptr[dev] )
	{
		/* set-up, once per device, with that device current */
		void *p = nullptr;
		CK( cudaGetSymbolAddress( &p, g_static_blit ));
		if( !p ) { snprintf( D.err, sizeof( D.err ), "rc_dev_blit_static: cudaGetSymbolAddress returned no address" ); return -1; }
		g_static_blit_ptr[dev] = reinterpret_cast<rc_static_blit_t *>( p );
	}
	rc_static_blit_t *sb = g_static_blit_ptr[dev];
	/* default stream, synchronous copies: no stream, event or buffer is created here */
	CK( cudaMemcpy( sb->lut, table, sizeof( uint32_t ) * 65536, cudaMemcpyHostToDevice ));
	CK( cudaMemcpy( sb->in, pixels, npx * sizeof( uint16_t ), cudaMemcpyHostToDevice ));
	const size_t want = ( npx / 8 + 255 ) / 256;
	const int grid = (int)( want < 148 * 8 ? ( want ? want : 1 ) : 148 * 8 );
	if( bpp == 2 ) k_blit<2><<<grid, 256>>>( sb->in, sb->lut, sb->out, npx );
	else k_blit<4><<<grid, 256>>>( sb->in, sb->lut, sb->out, npx );
	D.launches++;
	CK( cudaGetLastError());
	CK( cudaMemcpy( dst, sb->out, npx * (size_t)bpp, cudaMemcpyDeviceToHost ));
	return 0;
}

int rc_dev_read_counters( uint64_t *out, int n )
{
	unsigned long long h[RC_CTR_COUNT];
	CK( cudaMemcpy( h, D.d_counters, sizeof( h ), cudaMemcpyDeviceToHost ));
	for( int i = 0; i < n && i < RC_CTR_COUNT; i++ ) out[i] = h[i];
	return 0;
}

int rc_dev_read_view_surfs( int view, rc_surf_t *out, int max, int *n )
{
	if( !D.have_batch || view < 0 || view >= D.w.nviews ) return -1;
	int32_t nc;
	CK( cudaMemcpy( &nc, D.w.ncands + view, sizeof( nc ), cudaMemcpyDeviceToHost ));
	*n = nc;
	if( nc > max ) nc = max;
	if( nc > 0 ) CK( cudaMemcpy( out, D.w.surfs + (size_t)view * D.w.max_cands, sizeof( rc_surf_t ) * nc, cudaMemcpyDeviceToHost ));
	return 0;
}

int rc_dev_read_view_edges( int view, rc_edge_t *out, int max, int *n )
{
	if( !D.have_batch || view < 0 || view >= D.w.nviews ) return -1;
	int32_t ne;
	CK( cudaMemcpy( &ne, D.w.nedges + view, sizeof( ne ), cudaMemcpyDeviceToHost ));
	if( ne > D.w.max_edges ) ne = D.w.max_edges;
	*n = ne;
	if( ne > max ) ne = max;
	if( ne > 0 ) CK( cudaMemcpy( out, D.w.edges + (size_t)view * D.w.max_edges, sizeof( rc_edge_t ) * ne, cudaMemcpyDeviceToHost ));
	return 0;
}

} /* extern "C" */
