/*
 * rc_core.h - the per-item arithmetic of the ref_cuda frame path, written once as scalar
 * __host__ __device__ functions.  The CUDA kernels in rc_device.cu supply the parallel
 * decomposition (warp per view, thread per face, warp per scanline, thread per texel) around
 * these; tests/emu/ compiles the very same functions with g++ to debug integer parity against
 * the oracle on a box without a GPU (debug harness only - libref_cuda.so contains no CPU path).
 *
 * Float rule: every float expression keeps the reference's operand order and is compiled
 * without FMA contraction (nvcc -fmad=false, IEEE div/sqrt), matching the oracle's plain
 * gcc -O2 x86-64 SSE build (oracle/Makefile).  float->int uses rc_f2i (x86 cvttss2si
 * semantics incl. the 0x80000000 overflow result).
 */
#ifndef RC_CORE_H
#define RC_CORE_H

#include "rc_shared.h"
#include <stdlib.h>
#include <math.h>

#if defined( __CUDACC__ )
#define RC_HD __host__ __device__ __forceinline__
#define RC_HD_NOINLINE __host__ __device__ __noinline__
#else
#define RC_HD static inline
#define RC_HD_NOINLINE static
#endif

#define RC_NEAR_CLIP        0.01f  /* ref/soft/r_local.h:626 */
#define RC_BACKFACE_EPSILON 0.01   /* double constant, ref/soft/r_local.h */
#define RC_CYCLE            128    /* ref/soft/r_local.h:595 */
#define RC_BACKGROUND_KEY   0x7FFFFFFF

/* (int)f with x86-64 cvttss2si semantics: out-of-range and NaN give INT_MIN */
RC_HD int32_t rc_f2i( float f )
{
	if( !( f > -2147483904.0f && f < 2147483648.0f ))
		return (int32_t)0x80000000;
	return (int32_t)f;
}

RC_HD float rc_dot3( const float *a, const float *b )
{
	return a[0] * b[0] + a[1] * b[1] + a[2] * b[2];
}

/* TransformVector, ref/soft/r_misc.c:127 */
RC_HD void rc_transform( const rc_view_t *vw, const float *in, float *out )
{
	out[0] = rc_dot3( in, vw->vright );
	out[1] = rc_dot3( in, vw->vup );
	out[2] = rc_dot3( in, vw->vforward );
}

/* PlaneDiff, public/xash3d_mathlib.h:143 */
RC_HD float rc_planediff( const float *p, const rc_plane_t *pl )
{
	return ( pl->type < 3 ? p[pl->type] : rc_dot3( p, pl->normal )) - pl->dist;
}

/* ------------------------------------------------------------------------------------------ */
/* A1: view leaf, fat PVS, leaf->parent marking                                                 */

/* Mod_PointInLeaf, engine/common/mod_bmodel.c:1116: "<= 0" takes the back child */
RC_HD int32_t rc_point_in_leaf( const rc_scene_t *sc, const float *p )
{
	int32_t n = sc->headnode;
	while( n >= 0 )
	{
		const rc_node_t *node = &sc->nodes[n];
		n = node->child[rc_planediff( p, &sc->planes[node->plane] ) <= 0 ? 1 : 0];
	}
	return -1 - n;
}

/* Mod_DecompressPVS (mod_bmodel.c:1054) fused with Q_memor: OR one leaf's zero-run RLE row into out */
RC_HD void rc_pvs_or_row( const rc_scene_t *sc, int32_t visofs, uint8_t *out )
{
	int32_t n = sc->visbytes, d = 0;
	if( visofs < 0 )
	{
		for( d = 0; d < n; d++ ) out[d] = 0xFF;
		return;
	}
	const uint8_t *in = sc->visdata + visofs;
	while( d < n )
	{
		if( *in )
			out[d++] |= *in++;
		else
		{
			d += in[1];
			in += 2;
		}
	}
}

/* Mod_FatPVS (mod_bmodel.c:1203) with merge = fullvis = phs = false, as R_MarkLeaves calls it
 * (ref/soft/r_main.c:1030).  stack: scratch ints for the both-sides recursion. */
RC_HD void rc_fat_pvs( const rc_scene_t *sc, const float *org, float radius, uint8_t *out, int32_t *stack, int32_t stackcap )
{
	int32_t n = sc->visbytes;
	int32_t leaf = rc_point_in_leaf( sc, org );

	if( !sc->visdata_len || sc->leafs[leaf].cluster < 0 )
	{
		for( int32_t i = 0; i < n; i++ ) out[i] = 0xFF;
		return;
	}
	for( int32_t i = 0; i < n; i++ ) out[i] = 0;

	int32_t sp = 0;
	int32_t node = sc->headnode;
	for( ;; )
	{
		while( node >= 0 )
		{
			const rc_node_t *nd = &sc->nodes[node];
			float d = rc_planediff( org, &sc->planes[nd->plane] );
			if( d > radius ) node = nd->child[0];
			else if( d < -radius ) node = nd->child[1];
			else
			{
				/* go down both sides: front first, back deferred */
				if( sp < stackcap ) stack[sp++] = nd->child[1];
				node = nd->child[0];
			}
		}
		const rc_leaf_t *lf = &sc->leafs[-1 - node];
		if( lf->cluster >= 0 )
			rc_pvs_or_row( sc, lf->visofs, out );
		if( !sp ) break;
		node = stack[--sp];
	}
}

RC_HD int rc_testbit( const uint32_t *w, int32_t i ) { return ( w[i >> 5] >> ( i & 31 )) & 1; }
RC_HD void rc_setbit( uint32_t *w, int32_t i ) { w[i >> 5] |= 1u << ( i & 31 ); }

/* R_MarkLeaves, ref/soft/r_main.c:1021-1047: for every PVS bit stamp the leaf and its parent chain */
RC_HD void rc_mark_leaves( const rc_scene_t *sc, const uint8_t *vis, uint32_t *nodevis, uint32_t *leafmark )
{
	for( int32_t i = 0; i < sc->numclusters; i++ )
	{
		if( !( vis[i >> 3] & ( 1 << ( i & 7 ))))
			continue;
		rc_setbit( leafmark, i + 1 );
		int32_t node = sc->leafs[i + 1].parent;
		while( node >= 0 && !rc_testbit( nodevis, node ))
		{
			rc_setbit( nodevis, node );
			node = sc->nodes[node].parent;
		}
	}
}

/* ------------------------------------------------------------------------------------------ */
/* A19: texture encode - GL_UploadTexture + GL_BuildMipMap (ref/soft/r_image.c:214-283,125-207) per output texel.
 * The reference box-filters the RGBA image in place, level after level (each level rounds: ( a + b + c + d ) >> 2), and
 * encodes every level to pixel_t.  A texel of level j is therefore a fixed tree of averages over a 2^j x 2^j block of the
 * source - evaluated here on its own, so that every output texel of every level is independent (one thread each).      */

/* pixel_t code of an RGB triple already reduced to 5/6/5 bits: major RGB332 << 8 | minor GBRGBRGB (r_image.c:261-271) */
RC_HD uint32_t rc_pixel_encode( uint32_t r, uint32_t g, uint32_t b )
{
	const uint32_t major = ((( r >> 2 ) & 7u ) << 5 ) | ((( g >> 3 ) & 7u ) << 2 ) | (( b >> 3 ) & 3u );
	const uint32_t minor = ((( r >> 1 ) & 1u ) << 5 ) | (( r & 1u ) << 2 )
		| ((( g >> 2 ) & 1u ) << 7 ) | ((( g >> 1 ) & 1u ) << 4 ) | (( g & 1u ) << 1 )
		| ((( b >> 2 ) & 1u ) << 6 ) | ((( b >> 1 ) & 1u ) << 3 ) | ( b & 1u );
	return ( major << 8 | minor ) & 0xFFFFu;
}

/* one output texel of GL_BuildMipMap's box filter from the (up to) four texels of the level below it */
RC_HD uint32_t rc_tex_box( uint32_t a, uint32_t b, uint32_t c, uint32_t d, int two_x )
{
	uint32_t out = 0;
	if( two_x )
		for( int k = 0; k < 32; k += 8 )
			out |= (((( a >> k ) & 255u ) + (( b >> k ) & 255u ) + (( c >> k ) & 255u ) + (( d >> k ) & 255u )) >> 2 ) << k;
	else
		for( int k = 0; k < 32; k += 8 )
			out |= (((( a >> k ) & 255u ) + (( c >> k ) & 255u )) >> 1 ) << k;
	return out;
}

/* the RGBA value (packed, R in the low byte) GL_BuildMipMap leaves at (x, y) of a level; W, H: level-0 size.  A texel
 * averages its 2 x 2 block of the level below; a level of width (height) 1 has no second column (row): r_image.c:150-175.
 * One function per level (no recursion on the device). */
#define RC_TEX_LEVEL( NAME, BELOW, LEVEL ) \
RC_HD uint32_t NAME( const uint32_t *src, int32_t W, int32_t H, int32_t x, int32_t y ) \
{ \
	int32_t sw = W >> (( LEVEL ) - 1 ), sh = H >> (( LEVEL ) - 1 ); \
	if( sw < 1 ) sw = 1; \
	if( sh < 1 ) sh = 1; \
	const int two_x = ( 2 * x + 1 ) < sw, two_y = ( 2 * y + 1 ) < sh; \
	const uint32_t a = BELOW( src, W, H, 2 * x, 2 * y ); \
	const uint32_t c = two_y ? BELOW( src, W, H, 2 * x, 2 * y + 1 ) : a; \
	const uint32_t b = two_x ? BELOW( src, W, H, 2 * x + 1, 2 * y ) : 0u; \
	const uint32_t d = two_x ? ( two_y ? BELOW( src, W, H, 2 * x + 1, 2 * y + 1 ) : b ) : 0u; \
	return rc_tex_box( a, b, c, d, two_x ); \
}
RC_HD uint32_t rc_tex_rgba0( const uint32_t *src, int32_t W, int32_t H, int32_t x, int32_t y ) { (void)H; return src[(size_t)y * W + x]; }
RC_TEX_LEVEL( rc_tex_rgba1, rc_tex_rgba0, 1 )
RC_TEX_LEVEL( rc_tex_rgba2, rc_tex_rgba1, 2 )
RC_TEX_LEVEL( rc_tex_rgba3, rc_tex_rgba2, 3 )

RC_HD uint32_t rc_tex_rgba( const uint32_t *src, int32_t W, int32_t H, int32_t level, int32_t x, int32_t y )
{
	switch( level )
	{
	case 0: return rc_tex_rgba0( src, W, H, x, y );
	case 1: return rc_tex_rgba1( src, W, H, x, y );
	case 2: return rc_tex_rgba2( src, W, H, x, y );
	default: return rc_tex_rgba3( src, W, H, x, y );
	}
}

/* texel i of plane `plane` of a texture: 0..3 = the mip levels, 4 = the alpha plane (level 0: pixel >> 3 | alpha3 << 13) */
RC_HD uint16_t rc_tex_encode( const uint32_t *src, int32_t W, int32_t H, int32_t srcflags, int32_t plane, int32_t i )
{
	const int32_t level = plane == 4 ? 0 : plane;
	int32_t lw = W >> level;
	if( lw < 1 ) lw = 1;
	const int32_t y = i / lw, x = i - y * lw;
	const uint32_t v = rc_tex_rgba( src, W, H, level, x, y );
	const uint32_t r = ( v & 255u ) * 32u / 256u, g = (( v >> 8 ) & 255u ) * 64u / 256u, b = (( v >> 16 ) & 255u ) * 32u / 256u, a = v >> 24;
	const uint32_t px = rc_pixel_encode( r, g, b );
	if( plane == 4 )
		return (uint16_t)(( px >> 3 ) | (( a * 8u / 256u ) << 13 ));
	if( plane == 0 && ( srcflags & RC_TEXSRC_COLORKEY ) && a < 128u )
		return (uint16_t)0x0349; /* TRANSPARENT_COLOR */
	return (uint16_t)px;
}

/* ------------------------------------------------------------------------------------------ */
/* A2: front-to-back BSP walk                                                                   */

/* the 4-plane bbox test of R_RecursiveWorldNode (ref/soft/r_bsp.c:483-520); returns -1 when rejected,
 * else the reduced clipflags.  d is a double there, but ( (double)dotf - (double)dist ) has the sign of
 * the exact difference, so comparing the float dot against dist decides identically. */
RC_HD int32_t rc_cull_box( const rc_view_t *vw, const float *minmaxs, int32_t clipflags )
{
	for( int i = 0; i < 4; i++ )
	{
		if( !( clipflags & ( 1 << i )))
			continue;
		const int32_t *pindex = vw->frustum_idx[i];
		float pt[3];
		pt[0] = minmaxs[pindex[0]]; pt[1] = minmaxs[pindex[1]]; pt[2] = minmaxs[pindex[2]];
		double d = (double)rc_dot3( pt, vw->clip_normal[i] );
		d -= (double)vw->clip_dist[i];
		if( d <= 0 )
			return -1;
		pt[0] = minmaxs[pindex[3]]; pt[1] = minmaxs[pindex[4]]; pt[2] = minmaxs[pindex[5]];
		d = (double)rc_dot3( pt, vw->clip_normal[i] );
		d -= (double)vw->clip_dist[i];
		if( d >= 0 )
			clipflags &= ~( 1 << i );
	}
	return clipflags;
}

/* which side of a node plane the eye is on (r_bsp.c:550-571); float arithmetic, widened for the compares */
RC_HD double rc_node_dot( const rc_view_t *vw, const rc_plane_t *pl )
{
	float dotf;
	switch( pl->type )
	{
	case 0: dotf = vw->origin[0] - pl->dist; break;
	case 1: dotf = vw->origin[1] - pl->dist; break;
	case 2: dotf = vw->origin[2] - pl->dist; break;
	default: dotf = rc_dot3( vw->origin, pl->normal ) - pl->dist; break;
	}
	return (double)dotf;
}

/*
 * R_RenderWorld / R_RecursiveWorldNode (ref/soft/r_bsp.c:466-641) with the recursion unrolled onto an
 * explicit stack.  Output: the faces R_RenderFace would be called with, in call order (== ascending
 * surf_t.key order), each with its clipflags.  Keys: the reference bumps r_currentkey per posted face,
 * per node with surfaces and per leaf; only the order matters to the span generator, so the candidate
 * index is the key here (a strictly monotone relabelling).
 * stack entry = 2 ints: node ref, (clipflags | stage << 4).
 */
RC_HD int32_t rc_leaf_bmodel_slots( const rc_leaf_t *lf, const rc_bent_t *bents, int32_t nbents, int32_t upto );

/* bents / leafbase (optional): every visited leaf reserves the surface slots of the brush entities that touch it, at its
 * place in the order (= its key, LEAF_KEY r_bsp.c:541); leafbase[leaf] = the first of them (the caller presets -1) */
RC_HD int32_t rc_walk_world( const rc_scene_t *sc, const rc_view_t *vw, const uint32_t *nodevis, const uint32_t *leafmark,
	uint32_t *surfmark, int32_t *stack, int32_t stack_ints, rc_cand_t *cands, int32_t max_cands, int32_t *overflow,
	const rc_bent_t *bents, int32_t nbents, int32_t *leafbase, int32_t *leafkey )
{
	int32_t ncands = 0, sp = 0, nvisited = 0;
	stack[sp++] = sc->headnode; stack[sp++] = 15;

	while( sp )
	{
		int32_t st = stack[--sp];
		int32_t ref = stack[--sp];
		int32_t clipflags = st & 15;

		if( st & 16 )
		{
			/* second visit of a node: draw its faces, then descend the back side */
			const rc_node_t *node = &sc->nodes[ref];
			double dot = rc_node_dot( vw, &sc->planes[node->plane] );
			int side = dot >= 0 ? 0 : 1;
			int32_t c = node->numsurfaces;
			if( c )
			{
				int want;
				if( dot < -RC_BACKFACE_EPSILON ) want = 1;
				else if( dot > RC_BACKFACE_EPSILON ) want = 0;
				else want = -1;
				if( want >= 0 )
				{
					for( int32_t s = node->firstsurface; s < node->firstsurface + c; s++ )
					{
						int back = ( sc->surfaces[s].flags & RC_SURF_PLANEBACK ) ? 1 : 0;
						if( back == want && rc_testbit( surfmark, s ))
						{
							if( ncands < max_cands )
							{
								cands[ncands].surf = s;
								cands[ncands].clipflags = clipflags;
								ncands++;
							}
							else *overflow = 1;
						}
					}
				}
			}
			stack[sp++] = node->child[!side]; stack[sp++] = clipflags;
			continue;
		}

		if( ref < 0 )
		{
			int32_t li = -1 - ref;
			const rc_leaf_t *leaf = &sc->leafs[li];
			if( leaf->contents == RC_CONTENTS_SOLID || !rc_testbit( leafmark, li ))
				continue;
			if( clipflags && rc_cull_box( vw, leaf->minmaxs, clipflags ) < 0 )
				continue;
			for( int32_t m = leaf->firstmark; m < leaf->firstmark + leaf->nummark; m++ )
				rc_setbit( surfmark, sc->marksurfaces[m] );
			if( nbents && leafbase )
			{
				const int32_t n = rc_leaf_bmodel_slots( leaf, bents, nbents, nbents );
				leafbase[li] = ncands;
				if( leafkey ) leafkey[li] = nvisited++; /* the leaf's place in the walk: a key of its own even when it reserves nothing */
				for( int32_t k = 0; k < n; k++ )
				{
					if( ncands < max_cands ) { cands[ncands].surf = -1; cands[ncands].clipflags = 0; ncands++; }
					else *overflow = 1;
				}
			}
			continue;
		}

		const rc_node_t *node = &sc->nodes[ref];
		if( !rc_testbit( nodevis, ref ))
			continue;
		if( clipflags )
		{
			clipflags = rc_cull_box( vw, node->minmaxs, clipflags );
			if( clipflags < 0 )
				continue;
		}
		double dot = rc_node_dot( vw, &sc->planes[node->plane] );
		int side = dot >= 0 ? 0 : 1;
		if( sp + 4 > stack_ints ) { *overflow = 1; continue; }
		stack[sp++] = ref; stack[sp++] = clipflags | 16;
		stack[sp++] = node->child[side]; stack[sp++] = clipflags;
	}
	return ncands;
}

/* ------------------------------------------------------------------------------------------ */
/* A3: R_RenderFace / R_ClipEdge / R_EmitEdge (ref/soft/r_rast.c:63-584)                        */

/* derived per-surfedge tables (built once per map on the host): both vertices of every surfedge in winding order
 * (sv[numsurfedges][6]) and the other face that lists the same medge (partner[numsurfedges], see rc_scene_t) */
static inline void rc_build_sedge_tables( const rc_scene_t *sc, float *sv, int32_t *partner )
{
	for( int32_t i = 0; i < sc->numsurfedges; i++ )
	{
		partner[i] = -1;
		for( int k = 0; k < 6; k++ ) sv[(size_t)i * 6 + k] = 0.0f;
		int32_t l = sc->surfedges[i];
		const int a = l > 0 ? 0 : 1;
		if( l < 0 ) l = -l;
		if( l >= sc->numedges ) continue;
		const int32_t i0 = sc->edges[2 * l + a], i1 = sc->edges[2 * l + ( a ^ 1 )];
		if( i0 >= sc->numverts || i1 >= sc->numverts ) continue;
		for( int k = 0; k < 3; k++ )
		{
			sv[(size_t)i * 6 + k] = sc->verts[3 * i0 + k];
			sv[(size_t)i * 6 + 3 + k] = sc->verts[3 * i1 + k];
		}
	}
	/* first and second user of every medge, in face order */
	int32_t *user = (int32_t *)malloc( sizeof( int32_t ) * 2 * (size_t)( sc->numedges > 0 ? sc->numedges : 1 ));
	if( !user ) return;
	for( int32_t e = 0; e < 2 * sc->numedges; e++ ) user[e] = -1;
	for( int32_t s = 0; s < sc->numsurfaces; s++ )
		for( int32_t k = 0; k < sc->surfaces[s].numedges; k++ )
		{
			const int32_t se = sc->surfaces[s].firstedge + k;
			if( se < 0 || se >= sc->numsurfedges ) continue;
			int32_t l = sc->surfedges[se];
			if( l < 0 ) l = -l;
			if( l <= 0 || l >= sc->numedges ) continue;
			if( user[2 * l] < 0 ) user[2 * l] = se;
			else if( user[2 * l + 1] < 0 ) user[2 * l + 1] = se;
		}
	/* surfedge -> owning face */
	int32_t *owner = (int32_t *)malloc( sizeof( int32_t ) * (size_t)( sc->numsurfedges > 0 ? sc->numsurfedges : 1 ));
	if( owner )
	{
		for( int32_t i = 0; i < sc->numsurfedges; i++ ) owner[i] = -1;
		for( int32_t s = 0; s < sc->numsurfaces; s++ )
			for( int32_t k = 0; k < sc->surfaces[s].numedges; k++ )
			{
				const int32_t se = sc->surfaces[s].firstedge + k;
				if( se >= 0 && se < sc->numsurfedges ) owner[se] = s;
			}
		for( int32_t l = 1; l < sc->numedges; l++ )
		{
			const int32_t sa = user[2 * l], sb = user[2 * l + 1];
			if( sa < 0 || sb < 0 || owner[sa] < 0 || owner[sb] < 0 || owner[sa] == owner[sb] ) continue;
			const int32_t same = (( sc->surfedges[sa] > 0 ) == ( sc->surfedges[sb] > 0 )) ? 0x40000000 : 0;
			partner[sa] = owner[sb] | same;
			partner[sb] = owner[sa] | same;
		}
		free( owner );
	}
	free( user );
}

typedef struct
{
	/* statics of r_rast.c that live across the edges of one face */
	int     lastvertvalid, emitted, nearzionly;
	int     leftclipped, rightclipped;
	float   nearzi;
	float   u1, v1, lzi1;
	int32_t ceilv1;
	float   leftenter[3], leftexit[3], rightenter[3], rightexit[3];
	int     wrote; /* which of the four this face assigned: 1 leftexit, 2 leftenter, 4 rightexit, 8 rightenter */
	/* output */
	rc_edge_t *out;
	int32_t nout, maxout;
	uint32_t surfnum; /* candidate index + 1 */
} rc_facestate_t;

RC_HD void rc_project( const rc_view_t *vw, const float *world, float *pu, float *pv, float *plzi )
{
	float local[3], tv[3];
	local[0] = world[0] - vw->origin[0];
	local[1] = world[1] - vw->origin[1];
	local[2] = world[2] - vw->origin[2];
	rc_transform( vw, local, tv );
	if( tv[2] < RC_NEAR_CLIP )
		tv[2] = RC_NEAR_CLIP;
	float lzi = 1.0f / tv[2];
	float scale = vw->xscale * lzi;
	float u = ( vw->xcenter + scale * tv[0] );
	if( u < vw->fvrectx_adj ) u = vw->fvrectx_adj;
	if( u > vw->fvrectright_adj ) u = vw->fvrectright_adj;
	scale = vw->yscale * lzi;
	float v = ( vw->ycenter - scale * tv[1] );
	if( v < vw->fvrecty_adj ) v = vw->fvrecty_adj;
	if( v > vw->fvrectbottom_adj ) v = vw->fvrectbottom_adj;
	*pu = u; *pv = v; *plzi = lzi;
}

/* R_EmitEdge, r_rast.c:63-253.  The newedges[]/removeedges[] bucket insertion is replaced by the
 * flat per-view edge array; the scanline kernel sorts by u itself. */
RC_HD void rc_emit_edge( const rc_view_t *vw, rc_facestate_t *fs, const float *p0, const float *p1 )
{
	float u0, v0, lzi0;
	int32_t ceilv0;

	if( fs->lastvertvalid )
	{
		u0 = fs->u1; v0 = fs->v1; lzi0 = fs->lzi1; ceilv0 = fs->ceilv1;
	}
	else
	{
		rc_project( vw, p0, &u0, &v0, &lzi0 );
		ceilv0 = (int32_t)ceilf( v0 );
	}

	rc_project( vw, p1, &fs->u1, &fs->v1, &fs->lzi1 );

	if( fs->lzi1 > lzi0 ) lzi0 = fs->lzi1;
	if( lzi0 > fs->nearzi ) fs->nearzi = lzi0;

	if( fs->nearzionly )
		return;

	fs->emitted = 1;
	fs->ceilv1 = (int32_t)ceilf( fs->v1 );

	if( ceilv0 == fs->ceilv1 || ceilv0 < 0 )
		return; /* horizontal edge */

	int side = ceilv0 > fs->ceilv1;
	int32_t v, v2;
	float u, u_step;
	uint32_t surfs;

	if( side == 0 )
	{
		/* trailing edge (go from p1 to p2) */
		v = ceilv0; v2 = fs->ceilv1 - 1;
		surfs = fs->surfnum;
		u_step = (( fs->u1 - u0 ) / ( fs->v1 - v0 ));
		u = u0 + ((float)v - v0 ) * u_step;
	}
	else
	{
		/* leading edge (go from p2 to p1) */
		v2 = ceilv0 - 1; v = fs->ceilv1;
		surfs = fs->surfnum << 16;
		u_step = (( u0 - fs->u1 ) / ( v0 - fs->v1 ));
		u = fs->u1 + ((float)v - fs->v1 ) * u_step;
	}
	if( v < 0 || v >= 1200 ) /* MAXHEIGHT */
		return;

	int32_t iustep = rc_f2i( u_step * (float)0x100000 );
	int32_t iu = rc_f2i( u * (float)0x100000 + (float)0xFFFFF );
	if( iu < vw->vrect_x_adj_shift20 ) iu = vw->vrect_x_adj_shift20;
	if( iu > vw->vrectright_adj_shift20 ) iu = vw->vrectright_adj_shift20;

	if( fs->nout < fs->maxout )
	{
		rc_edge_t *e = &fs->out[fs->nout];
		e->u = iu; e->u_step = iustep;
		e->vv = (uint32_t)v | ((uint32_t)v2 << 16 );
		e->surfs = surfs;
	}
	fs->nout++;
}

/* R_ClipEdge, r_rast.c:261-352: the tail recursion over the clip-plane list becomes a loop */
RC_HD void rc_clip_edge( const rc_view_t *vw, rc_facestate_t *fs, const float *pv0, const float *pv1, const int *planes, int nplanes, int kstart )
{
	float a[3], b[3];
	a[0] = pv0[0]; a[1] = pv0[1]; a[2] = pv0[2];
	b[0] = pv1[0]; b[1] = pv1[1]; b[2] = pv1[2];

	for( int k = kstart; k < nplanes; k++ )
	{
		int pl = planes[k];
		float d0 = rc_dot3( a, vw->clip_normal[pl] ) - vw->clip_dist[pl];
		float d1 = rc_dot3( b, vw->clip_normal[pl] ) - vw->clip_dist[pl];

		if( d0 >= 0 )
		{
			if( d1 >= 0 )
				continue;
			/* only point 1 is clipped */
			float f = d0 / ( d0 - d1 );
			float c[3];
			c[0] = a[0] + f * ( b[0] - a[0] );
			c[1] = a[1] + f * ( b[1] - a[1] );
			c[2] = a[2] + f * ( b[2] - a[2] );
			if( pl == 0 ) { fs->leftclipped = 1; fs->wrote |= 1; fs->leftexit[0] = c[0]; fs->leftexit[1] = c[1]; fs->leftexit[2] = c[2]; }
			else if( pl == 1 ) { fs->rightclipped = 1; fs->wrote |= 4; fs->rightexit[0] = c[0]; fs->rightexit[1] = c[1]; fs->rightexit[2] = c[2]; }
			b[0] = c[0]; b[1] = c[1]; b[2] = c[2];
		}
		else
		{
			if( d1 < 0 )
				return; /* both points are clipped */
			/* only point 0 is clipped */
			fs->lastvertvalid = 0;
			float f = d0 / ( d0 - d1 );
			float c[3];
			c[0] = a[0] + f * ( b[0] - a[0] );
			c[1] = a[1] + f * ( b[1] - a[1] );
			c[2] = a[2] + f * ( b[2] - a[2] );
			if( pl == 0 ) { fs->leftclipped = 1; fs->wrote |= 2; fs->leftenter[0] = c[0]; fs->leftenter[1] = c[1]; fs->leftenter[2] = c[2]; }
			else if( pl == 1 ) { fs->rightclipped = 1; fs->wrote |= 8; fs->rightenter[0] = c[0]; fs->rightenter[1] = c[1]; fs->rightenter[2] = c[2]; }
			a[0] = c[0]; a[1] = c[1]; a[2] = c[2];
		}
	}
	rc_emit_edge( vw, fs, a, b );
}

/*
 * R_RenderFace's edge cache (medge16_t.cachededgeoffset, r_rast.c:447-478,139-160,315-318): of the two faces that share
 * an edge the one the walk posts first clips it and leaves one of three states behind; the second face then either
 * skips the edge altogether (cached as "fully clipped": really clipped away, or unclipped but horizontal), reuses the
 * emitted edge without clipping it again, or - the first face clipped it partially - processes it itself.  The state is
 * a pure function of the first face's clip planes and direction, so the second face recomputes it here.
 */
enum { RC_EDGE_OWN = 0, RC_EDGE_SKIP = 1, RC_EDGE_REUSE = 2 };

RC_HD int rc_first_face_edge_state( const rc_view_t *vw, const float *pv0, const float *pv1, int32_t clipflags )
{
	float a[3], b[3];
	a[0] = pv0[0]; a[1] = pv0[1]; a[2] = pv0[2];
	b[0] = pv1[0]; b[1] = pv1[1]; b[2] = pv1[2];
	int leftclipped = 0, clipped = 0;
	for( int pl = 0; pl < 4; pl++ )
	{
		if( !( clipflags & ( 1 << pl )))
			continue;
		float d0 = rc_dot3( a, vw->clip_normal[pl] ) - vw->clip_dist[pl];
		float d1 = rc_dot3( b, vw->clip_normal[pl] ) - vw->clip_dist[pl];
		if( d0 >= 0 )
		{
			if( d1 >= 0 )
				continue;
			float f = d0 / ( d0 - d1 );
			b[0] = a[0] + f * ( b[0] - a[0] ); b[1] = a[1] + f * ( b[1] - a[1] ); b[2] = a[2] + f * ( b[2] - a[2] );
		}
		else
		{
			if( d1 < 0 )
				return leftclipped ? RC_EDGE_OWN : RC_EDGE_SKIP; /* "if( !r_leftclipped ) cacheoffset = FULLY_CLIPPED_CACHED": even after a clip by another plane */
			float f = d0 / ( d0 - d1 );
			a[0] = a[0] + f * ( b[0] - a[0] ); a[1] = a[1] + f * ( b[1] - a[1] ); a[2] = a[2] + f * ( b[2] - a[2] );
		}
		clipped = 1;
		if( pl == 0 ) leftclipped = 1;
	}
	if( clipped )
		return RC_EDGE_OWN;
	/* emitted unclipped: a horizontal edge is cached as fully clipped, any other is reused */
	float u, v0, v1, lzi;
	rc_project( vw, a, &u, &v0, &lzi );
	rc_project( vw, b, &u, &v1, &lzi );
	return (int32_t)ceilf( v0 ) == (int32_t)ceilf( v1 ) ? RC_EDGE_SKIP : RC_EDGE_REUSE;
}

/* D_MipLevelForScale, ref/soft/r_edge.c:697-714 */
RC_HD int32_t rc_miplevel_for_scale( const rc_view_t *vw, float scale )
{
	int32_t lmiplevel;
	if( scale >= vw->scalemip[0] ) lmiplevel = 0;
	else if( scale >= vw->scalemip[1] ) lmiplevel = 1;
	else if( scale >= vw->scalemip[2] ) lmiplevel = 2;
	else lmiplevel = 3;
	if( lmiplevel < vw->minmip ) lmiplevel = vw->minmip;
	return lmiplevel;
}

/* D_CalcGradients, ref/soft/r_edge.c:741-813 (non-conveyor, non-TEX_WORLD_LUXELS faces) */
RC_HD void rc_calc_gradients( const rc_view_t *vw, const rc_surface_t *sf, int32_t miplevel, rc_surf_t *o )
{
	float p_saxis[3], p_taxis[3], p_temp1[3];
	float mipscale = 1.0f / (float)( 1 << miplevel );

	rc_transform( vw, sf->lmvecs[0], p_saxis );
	rc_transform( vw, sf->lmvecs[1], p_taxis );

	float t = vw->xscaleinv * mipscale;
	o->d_sdivzstepu = p_saxis[0] * t;
	o->d_tdivzstepu = p_taxis[0] * t;

	t = vw->yscaleinv * mipscale;
	o->d_sdivzstepv = -p_saxis[1] * t;
	o->d_tdivzstepv = -p_taxis[1] * t;

	o->d_sdivzorigin = p_saxis[2] * mipscale - vw->xcenter * o->d_sdivzstepu - vw->ycenter * o->d_sdivzstepv;
	o->d_tdivzorigin = p_taxis[2] * mipscale - vw->xcenter * o->d_tdivzstepu - vw->ycenter * o->d_tdivzstepv;

	p_temp1[0] = vw->modelorg_t[0] * mipscale;
	p_temp1[1] = vw->modelorg_t[1] * mipscale;
	p_temp1[2] = vw->modelorg_t[2] * mipscale;

	t = (float)0x10000 * mipscale;
	/* int - int + float: the sum is taken in float, then truncated on assignment (fixed16_t) */
	o->sadjust = rc_f2i((float)( rc_f2i( rc_dot3( p_temp1, p_saxis ) * (float)0x10000 + 0.5f )
		- (((int32_t)sf->lightmapmins[0] << 16 ) >> miplevel )) + sf->lmvecs[0][3] * t );
	o->tadjust = rc_f2i((float)( rc_f2i( rc_dot3( p_temp1, p_taxis ) * (float)0x10000 + 0.5f )
		- (((int32_t)sf->lightmapmins[1] << 16 ) >> miplevel )) + sf->lmvecs[1][3] * t );

	o->bbextents = (((int32_t)sf->lightextents[0] << 16 ) >> miplevel ) - 1;
	o->bbextentt = (((int32_t)sf->lightextents[1] << 16 ) >> miplevel ) - 1;
}

/* the edge loop of R_RenderFace for candidate ci: every surfedge clipped and emitted into fs (or skipped / reused as the
 * edge cache dictates); reports whether a left re-entry edge / a right nearzi pass is due */
RC_HD void rc_face_edges( const rc_scene_t *sc, const rc_view_t *vw, const rc_cand_t *cands, const uint16_t *candof, int32_t ci,
	rc_facestate_t *fs, int *makeleftedge, int *makerightedge )
{
	const rc_cand_t *cand = &cands[ci];
	const rc_surface_t *sf = &sc->surfaces[cand->surf];
	int planes[4], nplanes = 0;
	for( int i = 0; i < 4; i++ )
		if( cand->clipflags & ( 1 << i ))
			planes[nplanes++] = i;

	fs->emitted = 0; fs->nearzi = 0; fs->nearzionly = 0; fs->lastvertvalid = 0; fs->wrote = 0;
	fs->u1 = fs->v1 = fs->lzi1 = 0; fs->ceilv1 = 0;
	*makeleftedge = 0; *makerightedge = 0;

	/* both vertices of every surfedge were laid out per surfedge at upload: one load per edge, the next edge's
	 * issued before this one is clipped */
	const float *sv = sc->sedgeverts + (size_t)sf->firstedge * 6;
	float cur[6], nxt[6];
	for( int k = 0; k < 6; k++ ) cur[k] = sf->numedges > 0 ? sv[k] : 0.0f;
	for( int32_t i = 0; i < sf->numedges; i++ )
	{
		if( i + 1 < sf->numedges )
			for( int k = 0; k < 6; k++ ) nxt[k] = sv[( i + 1 ) * 6 + k];
		/* the face on the other side of this edge, when the walk posted it before this one, decided the edge's fate */
		int state = RC_EDGE_OWN;
		const int32_t partner = sc->sedgepartner ? sc->sedgepartner[sf->firstedge + i] : -1;
		if( partner >= 0 )
		{
			const int32_t pc = (int32_t)candof[partner & 0x3FFFFFFF] - 1;
			if( pc >= 0 && pc < ci )
				state = ( partner & 0x40000000 ) ? rc_first_face_edge_state( vw, cur, cur + 3, cands[pc].clipflags )
					: rc_first_face_edge_state( vw, cur + 3, cur, cands[pc].clipflags );
		}
		fs->leftclipped = fs->rightclipped = 0;
		if( state == RC_EDGE_OWN )
		{
			rc_clip_edge( vw, fs, cur, cur + 3, planes, nplanes, 0 );
			if( fs->leftclipped ) *makeleftedge = 1;
			if( fs->rightclipped ) *makerightedge = 1;
			fs->lastvertvalid = 1;
		}
		else
		{
			if( state == RC_EDGE_REUSE )
			{
				/* R_EmitCachedEdge: the edge as the first face emitted it - unclipped */
				fs->lastvertvalid = 0;
				rc_emit_edge( vw, fs, cur, cur + 3 );
			}
			fs->lastvertvalid = 0;
		}
		for( int k = 0; k < 6; k++ ) cur[k] = nxt[k];
	}
}

/* the two passes that follow the edge loop: the left re-entry edge and the right-side nearzi */
RC_HD void rc_face_side_edges( const rc_view_t *vw, const rc_cand_t *cand, rc_facestate_t *fs, int makeleftedge, int makerightedge )
{
	int planes[4], nplanes = 0;
	for( int i = 0; i < 4; i++ )
		if( cand->clipflags & ( 1 << i ))
			planes[nplanes++] = i;
	/* if there was a clip off the left edge, add that edge too (pclip->next: skip the left plane) */
	if( makeleftedge )
	{
		fs->lastvertvalid = 0;
		rc_clip_edge( vw, fs, fs->leftexit, fs->leftenter, planes, nplanes, 1 );
	}
	/* if there was a clip off the right edge, get the right r_nearzi (view_clipplanes[1].next) */
	if( makerightedge )
	{
		fs->lastvertvalid = 0;
		fs->nearzionly = 1;
		int k = 0;
		while( k < nplanes && planes[k] <= 1 ) k++;
		rc_clip_edge( vw, fs, fs->rightexit, fs->rightenter, planes, nplanes, k );
	}
}

/*
 * R_RenderFace (r_rast.c:384-586) for a world face + the per-surface part of D_SolidSurf /
 * D_TurbulentSurf / D_BackgroundSurf (r_edge.c:823-1053), in three steps:
 *   rc_face_begin   the edge loop (rc_face_edges); leaves a record of what the face did to r_leftexit / r_leftenter /
 *                   r_rightexit / r_rightenter, file statics of r_rast.c
 *   rc_face_inherit for the rare face whose left (right) plane clipped only one of the two edges that cross it - the
 *                   other was skipped through the edge cache - the missing point is whatever an EARLIER face of the frame
 *                   left in the static: the records of the faces before it are searched, last one first.  A point no
 *                   earlier face of the view wrote would come from the previous frame in the reference; here it is the origin.
 *   rc_face_finish  left re-entry edge, right-side nearzi, surf_t fields, mip level, gradients.
 * Edges go to fs->out (count in fs->nout, which may exceed maxout: the caller treats that as overflow).
 */
typedef struct
{
	float   pts[4][3];  /* leftexit, leftenter, rightexit, rightenter as the face left them */
	int32_t wrote;      /* bit i: pts[i] assigned by this face */
	int32_t state;      /* 1 make left edge, 2 make right edge, 4 emitted, 8 waits for rc_face_inherit */
	float   nearzi;
	int32_t nout;
} rc_facerec_t; /* 64 B */

RC_HD int rc_face_need( int makeleftedge, int makerightedge, int wrote )
{
	return ( makeleftedge ? 3 & ~wrote : 0 ) | ( makerightedge ? 12 & ~wrote : 0 );
}

RC_HD void rc_face_store( const rc_facestate_t *fs, int makeleftedge, int makerightedge, int waits, rc_facerec_t *r )
{
	for( int q = 0; q < 3; q++ )
	{
		r->pts[0][q] = fs->leftexit[q]; r->pts[1][q] = fs->leftenter[q];
		r->pts[2][q] = fs->rightexit[q]; r->pts[3][q] = fs->rightenter[q];
	}
	r->wrote = fs->wrote;
	r->state = ( makeleftedge ? 1 : 0 ) | ( makerightedge ? 2 : 0 ) | ( fs->emitted ? 4 : 0 ) | ( waits ? 8 : 0 );
	r->nearzi = fs->nearzi;
	r->nout = fs->nout;
}

RC_HD int rc_face_begin( const rc_scene_t *sc, const rc_view_t *vw, const rc_cand_t *cands, const uint16_t *candof, int32_t ci,
	rc_facestate_t *fs, int *makeleftedge, int *makerightedge )
{
	for( int k = 0; k < 3; k++ ) fs->leftexit[k] = fs->leftenter[k] = fs->rightexit[k] = fs->rightenter[k] = 0.0f;
	rc_face_edges( sc, vw, cands, candof, ci, fs, makeleftedge, makerightedge );
	return rc_face_need( *makeleftedge, *makerightedge, fs->wrote );
}

/* recs: the records of the view's candidates (those before ci are read) */
RC_HD void rc_face_inherit( const rc_facerec_t *recs, int32_t ci, int need, rc_facestate_t *fs )
{
	for( int32_t k = ci - 1; k >= 0 && need; k-- )
	{
		const int take = need & recs[k].wrote;
		if( !take )
			continue;
		if( take & 1 ) for( int q = 0; q < 3; q++ ) fs->leftexit[q] = recs[k].pts[0][q];
		if( take & 2 ) for( int q = 0; q < 3; q++ ) fs->leftenter[q] = recs[k].pts[1][q];
		if( take & 4 ) for( int q = 0; q < 3; q++ ) fs->rightexit[q] = recs[k].pts[2][q];
		if( take & 8 ) for( int q = 0; q < 3; q++ ) fs->rightenter[q] = recs[k].pts[3][q];
		need &= ~take;
	}
}

RC_HD void rc_face_finish( const rc_scene_t *sc, const rc_view_t *vw, const rc_cand_t *cand, rc_facestate_t *fs,
	int makeleftedge, int makerightedge, rc_surf_t *o )
{
	const rc_surface_t *sf = &sc->surfaces[cand->surf];
	rc_face_side_edges( vw, cand, fs, makeleftedge, makerightedge );

	o->surf = cand->surf;
	o->flags = sf->flags;
	o->emitted = fs->emitted;
	o->nearzi = fs->nearzi;
	if( !fs->emitted )
		return;

	const rc_plane_t *pl = &sc->planes[sf->plane];
	float p_normal[3];
	rc_transform( vw, pl->normal, p_normal );
	float distinv = 1.0f / ( pl->dist - rc_dot3( vw->origin, pl->normal ));
	o->d_zistepu = p_normal[0] * vw->xscaleinv * distinv;
	o->d_zistepv = -p_normal[1] * vw->yscaleinv * distinv;
	o->d_ziorigin = p_normal[2] * distinv - vw->xcenter * o->d_zistepu - vw->ycenter * o->d_zistepv;

	if( sf->flags & RC_SURF_DRAWSKY )
	{
		/* D_BackgroundSurf: flat fill at "infinity" */
		o->d_zistepu = 0; o->d_zistepv = 0; o->d_ziorigin = -0.9f;
		o->miplevel = 0; o->cache_ofs = -1; o->cachewidth = 0;
		return;
	}
	if( sf->flags & RC_SURF_DRAWTURB )
	{
		/* D_TurbulentSurf: mip 0 of the 64x64 texture itself, no cache */
		o->miplevel = 0;
		o->cache_ofs = sc->textures[sf->texnum].ofs[0];
		o->cachewidth = 64;
		rc_calc_gradients( vw, sf, 0, o );
		return;
	}

	/* D_AlphaSurf (r_edge.c:931) always draws from mip 0; D_SolidSurf picks the level from the nearest 1/z */
	int32_t miplevel = ( cand->clipflags & 0x100 ) ? 0 : rc_miplevel_for_scale( vw, fs->nearzi * vw->scale_for_mip );
	while(( 1 << miplevel ) > sf->sample_size )
		miplevel--;
	o->miplevel = miplevel;
	o->cache_ofs = sf->cache_ofs[miplevel];
	o->cachewidth = (int32_t)sf->lightextents[0] >> miplevel;
	rc_calc_gradients( vw, sf, miplevel, o );
}

/* ------------------------------------------------------------------------------------------ */
/* A13: opaque brush entities - R_DrawBEntitiesOnList (ref/soft/r_main.c:781), R_RotateBmodel,  */
/* R_RecursiveClipBPoly, R_DrawSolidClippedSubmodelPolygons, R_DrawSubmodelPolygons             */
/* (r_bsp.c:97-451), R_RenderBmodelFace (r_rast.c:596)                                          */

#if defined( __CUDA_ARCH__ )
#define RC_FETCH_ADD( p, n ) atomicAdd(( p ), ( n ))
#define RC_ATOMIC_OR( p, v ) atomicOr(( p ), ( v ))
#else
#define RC_ATOMIC_OR( p, v ) ( *( p ) |= ( v ))
static inline int32_t rc_fetch_add_host( int32_t *p, int32_t n ) { const int32_t o = *p; *p += n; return o; }
#define RC_FETCH_ADD( p, n ) rc_fetch_add_host(( p ), ( n ))
#endif

#define RC_TOPNODE_NONE    ((int32_t)0x80000000 )
#define RC_BM_MAXVERTS 56  /* per face: its own vertices + the points the BSP planes cut (the reference: 1000 per face) */
#define RC_BM_MAXEDGES 112
#define RC_BM_STACK    48
#define RC_BM_FRAGEDGES 24 /* edges one fragment can emit: its polygon's + the left re-entry edge */

/* R_BmodelCheckBBox, r_main.c:640: -1 = BMODEL_FULLY_CLIPPED, else the planes the box crosses */
RC_HD int32_t rc_bmodel_check_bbox( const rc_view_t *vw, const float *minmaxs )
{
	int32_t clipflags = 0;
	for( int i = 0; i < 4; i++ )
	{
		const int32_t *pindex = vw->frustum_idx[i];
		float pt[3];
		pt[0] = minmaxs[pindex[0]]; pt[1] = minmaxs[pindex[1]]; pt[2] = minmaxs[pindex[2]];
		float d = rc_dot3( pt, vw->clip_normal[i] );
		d -= vw->clip_dist[i];
		if( d <= 0 )
			return -1;
		pt[0] = minmaxs[pindex[3]]; pt[1] = minmaxs[pindex[4]]; pt[2] = minmaxs[pindex[5]];
		d = rc_dot3( pt, vw->clip_normal[i] );
		d -= vw->clip_dist[i];
		if( d <= 0 )
			clipflags |= 1 << i;
	}
	return clipflags;
}

/* BOX_ON_PLANE_SIDE / BoxOnPlaneSide, public/xash3d_mathlib.h:197, xash3d_mathlib.c:368 */
RC_HD int rc_box_on_plane_side( const float *emins, const float *emaxs, const rc_plane_t *p )
{
	if( p->type < 3 )
	{
		if( p->dist <= emins[p->type] ) return 1;
		if( p->dist >= emaxs[p->type] ) return 2;
		return 3;
	}
	/* dist1: the corner furthest along the normal, dist2: the opposite one (the signbits switch) */
	float a[3], b[3];
	for( int k = 0; k < 3; k++ )
	{
		const int neg = p->normal[k] < 0;
		a[k] = neg ? emins[k] : emaxs[k];
		b[k] = neg ? emaxs[k] : emins[k];
	}
	const float dist1 = p->normal[0] * a[0] + p->normal[1] * a[1] + p->normal[2] * a[2];
	const float dist2 = p->normal[0] * b[0] + p->normal[1] * b[1] + p->normal[2] * b[2];
	int sides = 0;
	if( dist1 >= p->dist ) sides = 1;
	if( dist2 < p->dist ) sides |= 2;
	return sides;
}

/* R_FindTopnode, r_main.c:684: a node index, -1 - leaf for a non-solid leaf, RC_TOPNODE_NONE for "not visible" */
RC_HD int32_t rc_find_topnode( const rc_scene_t *sc, const uint32_t *nodevis, const uint32_t *leafvis, const float *minmaxs )
{
	int32_t node = sc->headnode;
	for( ;; )
	{
		if( node < 0 )
		{
			const int32_t leaf = -1 - node;
			if( !rc_testbit( leafvis, leaf ))
				return RC_TOPNODE_NONE;
			return sc->leafs[leaf].contents != -2 /* CONTENTS_SOLID */ ? node : RC_TOPNODE_NONE;
		}
		if( !rc_testbit( nodevis, node ))
			return RC_TOPNODE_NONE;
		const int sides = rc_box_on_plane_side( minmaxs, minmaxs + 3, &sc->planes[sc->nodes[node].plane] );
		if( sides == 3 )
			return node;
		node = ( sides & 1 ) ? sc->nodes[node].child[0] : sc->nodes[node].child[1];
	}
}

/* R_EntityRotate, r_bsp.c:81 */
RC_HD void rc_entity_rotate( const rc_bent_t *be, const float *in, float *out )
{
	out[0] = rc_dot3( be->rot[0], in );
	out[1] = rc_dot3( be->rot[1], in );
	out[2] = rc_dot3( be->rot[2], in );
}

/* the view as a brush entity's faces see it: modelorg = vieworigin - r_entorigin, then R_RotateBmodel (modelorg and the view
 * vectors through entity_rotation, R_TransformFrustum again).  transformed_modelorg is D_SolidSurf's (r_edge.c:1003): the
 * unrotated difference through the BASE view vectors. */
RC_HD void rc_bmodel_view( const rc_view_t *vw, const rc_bent_t *be, rc_view_t *bv )
{
	*bv = *vw;
	float mo[3];
	mo[0] = vw->origin[0] - be->origin[0]; mo[1] = vw->origin[1] - be->origin[1]; mo[2] = vw->origin[2] - be->origin[2];
	rc_transform( vw, mo, bv->modelorg_t );
	rc_entity_rotate( be, mo, bv->origin );
	rc_entity_rotate( be, vw->vforward, bv->vforward );
	rc_entity_rotate( be, vw->vright, bv->vright );
	rc_entity_rotate( be, vw->vup, bv->vup );
	for( int i = 0; i < 4; i++ )
	{
		float v[3], v2[3];
		v[0] = vw->screenedge[i][2]; v[1] = -vw->screenedge[i][0]; v[2] = vw->screenedge[i][1];
		v2[0] = v[1] * bv->vright[0] + v[2] * bv->vup[0] + v[0] * bv->vforward[0];
		v2[1] = v[1] * bv->vright[1] + v[2] * bv->vup[1] + v[0] * bv->vforward[1];
		v2[2] = v[1] * bv->vright[2] + v[2] * bv->vup[2] + v[0] * bv->vforward[2];
		bv->clip_normal[i][0] = v2[0]; bv->clip_normal[i][1] = v2[1]; bv->clip_normal[i][2] = v2[2];
		bv->clip_dist[i] = rc_dot3( bv->origin, v2 );
	}
}

/* does the entity's box touch the leaf's?  k_viswalk reserves surface slots (= the leaf's key) for every entity that does */
RC_HD int rc_bent_touches_leaf( const rc_bent_t *be, const rc_leaf_t *lf )
{
	for( int k = 0; k < 3; k++ )
		if( be->minmaxs[k] > lf->minmaxs[3 + k] || be->minmaxs[3 + k] < lf->minmaxs[k] )
			return 0;
	return 1;
}

/* surface slots a visited leaf needs: the faces of every brush entity whose box touches it */
RC_HD int32_t rc_leaf_bmodel_slots( const rc_leaf_t *lf, const rc_bent_t *bents, int32_t nbents, int32_t upto )
{
	int32_t n = 0;
	for( int32_t e = 0; e < nbents && e < upto; e++ )
		if( !bents[e].trans && rc_bent_touches_leaf( &bents[e], lf ))
			n += bents[e].numfaces;
	return n;
}

typedef struct { int16_t v0, v1, next; } rc_bedge_t;

/* where a brush entity's fragments go: the view's surface and edge arrays */
typedef struct
{
	rc_surf_t *surfs; int32_t max_cands;
	rc_edge_t *edges; int32_t max_edges; int32_t *nedges;
	const int32_t *leafbase;  /* [numleafs] first reserved slot of every visited leaf, -1 = not visited */
	const uint16_t *candof;
	const int32_t *leafkey;   /* NULL: the pieces go into the leaf's reserved slots of the view's surface list (opaque entities).
	                           * else [numleafs] place of every visited leaf in the walk: the entity keeps a surface list of its
	                           * own (R_DrawBrushModel's edge frame), pieces in arrival order, keyed by leafkey + 1, mip 0 */
	uint32_t *cache_need;     /* (face, mip) blocks the batch samples (device); NULL: the caller derives them from the surfaces */
	int32_t nsurfs, overflow, noutmax;
} rc_bmout_t;

/* post one fragment (or whole face): surface record into its slot, edges appended to the view's list */
RC_HD void rc_bmodel_post( const rc_scene_t *sc, const rc_view_t *bv, const rc_cand_t *cand, rc_facestate_t *fs, int mkleft, int mkright,
	int32_t slot, int32_t subkey, rc_bmout_t *out )
{
	rc_surf_t o;
	rc_face_finish( sc, bv, cand, fs, mkleft, mkright, &o );
	if( !o.emitted )
		return;
	if( fs->nout > fs->maxout ) { out->overflow = 1; return; }
	o.emitted = 1 | ( subkey << 8 );
	const int32_t at = RC_FETCH_ADD( out->nedges, fs->nout );
	if( at + fs->nout > out->noutmax ) out->noutmax = at + fs->nout;
	if( at + fs->nout <= out->max_edges )
		for( int32_t k = 0; k < fs->nout; k++ ) out->edges[at + k] = fs->out[k];
	out->surfs[slot] = o;
	out->nsurfs++;
	if( out->cache_need && !( o.flags & ( RC_SURF_DRAWSKY | RC_SURF_DRAWTURB )) && o.cache_ofs >= 0 )
	{
		const int32_t cs = cand->surf * 4 + o.miplevel;
		RC_ATOMIC_OR( &out->cache_need[cs >> 5], 1u << ( cs & 31 ));
	}
}

RC_HD void rc_bm_fs_init( rc_facestate_t *fs, rc_edge_t *buf, int32_t slot )
{
	fs->emitted = 0; fs->nearzi = 0; fs->nearzionly = 0; fs->lastvertvalid = 0; fs->wrote = 0;
	fs->leftclipped = fs->rightclipped = 0;
	fs->u1 = fs->v1 = fs->lzi1 = 0; fs->ceilv1 = 0;
	for( int k = 0; k < 3; k++ ) fs->leftexit[k] = fs->leftenter[k] = fs->rightexit[k] = fs->rightenter[k] = 0.0f;
	fs->out = buf; fs->nout = 0; fs->maxout = RC_BM_FRAGEDGES; fs->surfnum = (uint32_t)( slot + 1 );
}

/*
 * One face of one brush entity in one view.  topnode: rc_find_topnode's answer for the entity.
 *   a leaf:  R_DrawSubmodelPolygons - the face goes through R_RenderFace as a world face does (no edge cache: insubmodel)
 *   a node:  R_DrawSolidClippedSubmodelPolygons - the face's edge loop is cut by the BSP planes below topnode
 *            (R_RecursiveClipBPoly); every piece that ends in a visible non-solid leaf is posted (R_RenderBmodelFace)
 * All pieces in one leaf carry that leaf's key (subkey = its first reserved slot + 1): the span stage sorts them among the
 * world's surfaces by it, and among themselves by 1/z.
 */
RC_HD void rc_bmodel_face( const rc_scene_t *sc, const rc_view_t *bv, const rc_bent_t *bents, int32_t nbents, int32_t ent, int32_t face,
	int32_t clipflags, int32_t topnode, const uint32_t *nodevis, const uint32_t *leafvis, rc_bmout_t *out )
{
	const rc_bent_t *be = &bents[ent];
	const rc_surface_t *sf = &sc->surfaces[face];
	const rc_plane_t *pl = &sc->planes[sf->plane];
	const float dot = rc_dot3( bv->origin, pl->normal ) - pl->dist;
	const int back = ( sf->flags & RC_SURF_PLANEBACK ) != 0;
	rc_edge_t ebuf[RC_BM_FRAGEDGES];
	rc_facestate_t fs;
	rc_cand_t cand;
	cand.surf = face; cand.clipflags = clipflags | ( out->leafkey ? 0x100 : 0 ); /* 0x100: D_AlphaSurf draws from mip 0 */

	if( topnode < 0 )
	{
		if( !(( back && dot < -RC_BACKFACE_EPSILON ) || ( !back && dot > RC_BACKFACE_EPSILON )))
			return;
		const int32_t leaf = -1 - topnode;
		const int32_t base = out->leafbase[leaf];
		if( base < 0 )
			return; /* the walk never reached the leaf (outside the frustum): nothing of it can be on screen */
		const int32_t slot = out->leafkey ? out->nsurfs : base + rc_leaf_bmodel_slots( &sc->leafs[leaf], bents, nbents, ent ) + ( face - be->firstface );
		if( slot >= out->max_cands ) { out->overflow = 1; return; }
		int mkl, mkr;
		rc_bm_fs_init( &fs, ebuf, slot );
		rc_face_edges( sc, bv, &cand, out->candof, 0, &fs, &mkl, &mkr );
		rc_bmodel_post( sc, bv, &cand, &fs, mkl, mkr, slot, out->leafkey ? out->leafkey[leaf] + 1 : base + 1, out );
		return;
	}

	if(( !back && dot < -RC_BACKFACE_EPSILON ) || ( back && dot > RC_BACKFACE_EPSILON ))
		return;

	float verts[RC_BM_MAXVERTS][3];
	rc_bedge_t edges[RC_BM_MAXEDGES];
	int32_t stk_node[RC_BM_STACK]; int16_t stk_head[RC_BM_STACK];
	int32_t numbverts = 0, numbedges = 0, sp = 0;
	int16_t pfrontenter = 0, pfrontexit = 0;

	/* the face's edge loop, clockwise: vertex i -> i + 1 */
	if( sf->numedges > RC_BM_MAXVERTS / 2 || sf->numedges < 1 ) { out->overflow = 1; return; }
	const float *sv = sc->sedgeverts + (size_t)sf->firstedge * 6;
	for( int32_t j = 0; j < sf->numedges; j++ )
	{
		verts[numbverts][0] = sv[j * 6]; verts[numbverts][1] = sv[j * 6 + 1]; verts[numbverts][2] = sv[j * 6 + 2];
		edges[j].v0 = (int16_t)numbverts;
		edges[j].v1 = (int16_t)(( j + 1 < sf->numedges ) ? numbverts + 1 : 0 );
		edges[j].next = (int16_t)(( j + 1 < sf->numedges ) ? j + 1 : -1 );
		numbverts++;
	}
	numbedges = sf->numedges;
	stk_node[0] = topnode; stk_head[0] = 0; sp = 1;

	int planes[4], nplanes = 0;
	for( int i = 0; i < 4; i++ )
		if( clipflags & ( 1 << i ))
			planes[nplanes++] = i;

	while( sp > 0 )
	{
		sp--;
		const int32_t node = stk_node[sp];
		int16_t pedges = stk_head[sp];
		if( node < 0 )
		{
			/* a leaf: draw if it is visible and not solid */
			const int32_t leaf = -1 - node;
			if( !rc_testbit( leafvis, leaf ) || sc->leafs[leaf].contents == -2 )
				continue;
			const int32_t base = out->leafbase[leaf];
			if( base < 0 )
				continue;
			const int32_t slot = out->leafkey ? out->nsurfs : base + rc_leaf_bmodel_slots( &sc->leafs[leaf], bents, nbents, ent ) + ( face - be->firstface );
			if( slot >= out->max_cands ) { out->overflow = 1; continue; }
			/* R_RenderBmodelFace */
			int mkl = 0, mkr = 0;
			rc_bm_fs_init( &fs, ebuf, slot );
			for( ; pedges >= 0; pedges = edges[pedges].next )
			{
				fs.leftclipped = fs.rightclipped = 0;
				rc_clip_edge( bv, &fs, verts[edges[pedges].v0], verts[edges[pedges].v1], planes, nplanes, 0 );
				if( fs.leftclipped ) mkl = 1;
				if( fs.rightclipped ) mkr = 1;
			}
			rc_bmodel_post( sc, bv, &cand, &fs, mkl, mkr, slot, out->leafkey ? out->leafkey[leaf] + 1 : base + 1, out );
			continue;
		}
		if( !rc_testbit( nodevis, node ))
			continue;

		/* R_RecursiveClipBPoly: the BSP plane in model space */
		const rc_plane_t *sp_ = &sc->planes[sc->nodes[node].plane];
		float tn[3];
		const float tdist = sp_->dist - rc_dot3( be->origin, sp_->normal );
		tn[0] = rc_dot3( be->rot[0], sp_->normal ); tn[1] = rc_dot3( be->rot[1], sp_->normal ); tn[2] = rc_dot3( be->rot[2], sp_->normal );
		int16_t side_head[2] = { -1, -1 };
		int makeclippededge = 0, abandoned = 0;
		while( pedges >= 0 )
		{
			const int16_t pnext = edges[pedges].next;
			const int16_t iv0 = edges[pedges].v0, iv1 = edges[pedges].v1;
			const float lastdist = rc_dot3( verts[iv0], tn ) - tdist;
			const int lastside = lastdist > 0 ? 0 : 1;
			const float dist = rc_dot3( verts[iv1], tn ) - tdist;
			const int side = dist > 0 ? 0 : 1;
			if( side != lastside )
			{
				if( numbverts >= RC_BM_MAXVERTS || numbedges >= RC_BM_MAXEDGES - 1 ) { out->overflow = 1; abandoned = 1; break; }
				const float frac = lastdist / ( lastdist - dist );
				const int16_t pt = (int16_t)numbverts++;
				verts[pt][0] = verts[iv0][0] + frac * ( verts[iv1][0] - verts[iv0][0] );
				verts[pt][1] = verts[iv0][1] + frac * ( verts[iv1][1] - verts[iv0][1] );
				verts[pt][2] = verts[iv0][2] + frac * ( verts[iv1][2] - verts[iv0][2] );
				rc_bedge_t *e1 = &edges[numbedges];
				e1->next = side_head[lastside]; side_head[lastside] = (int16_t)numbedges; e1->v0 = iv0; e1->v1 = pt;
				rc_bedge_t *e2 = &edges[numbedges + 1];
				e2->next = side_head[side]; side_head[side] = (int16_t)( numbedges + 1 ); e2->v0 = pt; e2->v1 = iv1;
				numbedges += 2;
				if( side == 0 ) pfrontenter = pt; else pfrontexit = pt;
				makeclippededge = 1;
			}
			else
			{
				edges[pedges].next = side_head[side];
				side_head[side] = pedges;
			}
			pedges = pnext;
		}
		if( abandoned )
			continue;
		if( makeclippededge )
		{
			if( numbedges >= RC_BM_MAXEDGES - 2 ) { out->overflow = 1; continue; }
			rc_bedge_t *e1 = &edges[numbedges];
			e1->next = side_head[0]; side_head[0] = (int16_t)numbedges; e1->v0 = pfrontexit; e1->v1 = pfrontenter;
			rc_bedge_t *e2 = &edges[numbedges + 1];
			e2->next = side_head[1]; side_head[1] = (int16_t)( numbedges + 1 ); e2->v0 = pfrontenter; e2->v1 = pfrontexit;
			numbedges += 2;
		}
		/* side 0 is handled before side 1: push it last */
		for( int i = 1; i >= 0; i-- )
		{
			if( side_head[i] < 0 )
				continue;
			if( sp >= RC_BM_STACK ) { out->overflow = 1; continue; }
			stk_node[sp] = sc->nodes[node].child[i]; stk_head[sp] = side_head[i]; sp++;
		}
	}
}

/* ------------------------------------------------------------------------------------------ */
/* A4: span generation for one scanline (ref/soft/r_edge.c:360-583)                             */

typedef struct { int32_t u; uint32_t surfs; } rc_event_t; /* an active edge at this scanline */
typedef struct { int32_t surf; int16_t u, count; } rc_span_t;  /* 8 B; surf: 0 = background, else candidate index + 1 */

/* edge->u after (y - v) executions of "edge->u += edge->u_step" (R_StepActiveU, r_edge.c:187): int32 wraparound */
RC_HD int32_t rc_edge_u_at( const rc_edge_t *e, int32_t y )
{
	uint32_t v = e->vv & 0xFFFF;
	return (int32_t)((uint32_t)e->u + ((uint32_t)y - v ) * (uint32_t)e->u_step );
}

RC_HD int rc_edge_active( const rc_edge_t *e, int32_t y )
{
	int32_t v = (int32_t)( e->vv & 0xFFFF ), v2 = (int32_t)( e->vv >> 16 );
	return y >= v && y <= v2;
}

/* total order the active edge table is kept in: by u; ties (which only arise for edges clamped to a
 * screen border or for the two copies of a shared edge) broken so that the result is deterministic */
RC_HD uint64_t rc_event_key( int32_t u, uint32_t surfs )
{
	/* low word: leading surface in bits 0..15, trailing surface in bits 16..31 (the halves of rc_edge_t.surfs
	 * swapped) - at equal u a leading edge sorts in front of a trailing edge, the order R_EmitEdge's sorted insert
	 * gives new edges (u_check++ for trailing edges, ref/soft/r_rast.c:233-235) */
	return ((uint64_t)((uint32_t)u ^ 0x80000000u ) << 32 ) | ( surfs >> 16 ) | ( surfs << 16 );
}

/*
 * R_GenerateSpans + R_LeadingEdge + R_TrailingEdge + R_CleanupSpan for one scanline, events sorted by u.
 * The key-ordered surf_t stack is a sorted array (world keys are unique and equal the candidate
 * order); spanstate (r_edge.c:405,366) is kept as a count per stack entry plus a short list of
 * "inverted" surfaces whose trailing edge came first.
 * stk/cnt: scratch of stkcap entries; neg: scratch of 2*negcap ints.
 * Returns the number of spans written (spans of surf 0 = background are written too).
 */
/* the key R_LeadingEdge compares: a world surface's own number; a brush entity's piece carries its leaf's (all pieces in a
 * leaf share it: "all bmodels in a leaf share the same key", r_bsp.c:542) - the number of the leaf's first reserved slot */
RC_HD int32_t rc_span_key( const rc_surf_t *surfs, int32_t s )
{
	const int32_t sub = surfs ? ( surfs[s - 1].emitted >> 8 ) : 0;
	return sub ? sub : s;
}

/* "must be two bmodels in the same leaf; sort on 1/z" (R_LeadingEdge, r_edge.c:419-445,463-481): is the entering piece a
 * in front of the active piece b at this edge? */
RC_HD int rc_span_zi_in_front( const rc_surf_t *a, const rc_surf_t *b, int32_t u, int32_t y )
{
	const float fv = (float)y;
	const float fu = (float)( u - 0xFFFFF ) * ( 1.0f / 0x100000 );
	const float newzi = a->d_ziorigin + fv * a->d_zistepv + fu * a->d_zistepu;
	const float newzibottom = newzi * 0.99f;
	const float testzi = b->d_ziorigin + fv * b->d_zistepv + fu * b->d_zistepu;
	if( newzibottom >= testzi )
		return 1;
	const float newzitop = newzi * 1.01f;
	if( newzitop >= testzi && a->d_zistepu >= b->d_zistepu )
		return 1;
	return 0;
}

/* surfs / y: the view's surface records and the scanline - needed when brush entities are in the view (pieces that share
 * a leaf key are ordered by 1/z where they meet); NULL when there are none */
RC_HD int32_t rc_generate_spans( const unsigned long long *keys, int32_t nev, int32_t head_iu, int32_t tail_u, const uint16_t *bsurf,
	int32_t *stk, int32_t *cnt, int32_t stkcap, int32_t *neg, int32_t negcap, rc_span_t *spans, int32_t maxspans, int32_t *overflow,
	const rc_surf_t *surfs, int32_t y )
{
	int32_t depth = 0, nneg = 0, nspans = 0;
	int32_t top = 0, top_last_u = head_iu; /* surfaces[1].last_u = edge_head_u_shift20 */

#define RC_EMIT_SPAN( SURF, LAST, IU ) \
	if(( IU ) > ( LAST )) { if( nspans < maxspans ) { spans[nspans].u = (int16_t)( LAST ); spans[nspans].count = (int16_t)(( IU ) - ( LAST )); spans[nspans].surf = ( SURF ); nspans++; } else *overflow = 1; }

	for( int32_t e = 0; e < nev; e++ )
	{
		/* keys: the line's events as sorted rc_event_key values; bsurf (optional): band-local -> view surface numbers */
		const unsigned long long key = keys[e];
		const int32_t u = (int32_t)((uint32_t)( key >> 32 ) ^ 0x80000000u );
		if( u > tail_u )
			break; /* sorted past edge_tail: R_GenerateSpans stops there */
		int32_t iu = u >> 20;
		int32_t trail = (int32_t)((uint32_t)key >> 16 ), lead = (int32_t)((uint32_t)key & 0xFFFFu );
		if( bsurf ) { trail = bsurf[trail]; lead = bsurf[lead]; }

		if( trail )
		{
			/* R_TrailingEdge: --spanstate == 0 -> leave the stack */
			int32_t i;
			for( i = 0; i < depth; i++ ) if( stk[i] == trail ) break;
			if( i < depth )
			{
				if( --cnt[i] == 0 )
				{
					if( i == 0 )
					{
						RC_EMIT_SPAN( trail, top_last_u, iu )
						top_last_u = iu; /* surf->next->last_u = iu */
					}
					for( int32_t k = i; k + 1 < depth; k++ ) { stk[k] = stk[k + 1]; cnt[k] = cnt[k + 1]; }
					depth--;
					top = depth ? stk[0] : 0;
				}
			}
			else
			{
				/* inverted span: spanstate goes negative */
				for( i = 0; i < nneg; i++ ) if( neg[2 * i] == trail ) break;
				if( i < nneg ) neg[2 * i + 1]--;
				else if( nneg < negcap ) { neg[2 * nneg] = trail; neg[2 * nneg + 1] = -1; nneg++; }
				else *overflow = 1;
			}
		}
		if( lead )
		{
			/* R_LeadingEdge: ++spanstate == 1 -> enter the stack at its key position */
			int32_t i;
			for( i = 0; i < nneg; i++ ) if( neg[2 * i] == lead ) break;
			if( i < nneg )
			{
				if( ++neg[2 * i + 1] == 0 ) { neg[2 * i] = neg[2 * ( nneg - 1 )]; neg[2 * i + 1] = neg[2 * ( nneg - 1 ) + 1]; nneg--; }
				continue;
			}
			if( !surfs )
			{
				for( i = 0; i < depth; i++ ) if( stk[i] >= lead ) break;
				if( i < depth && stk[i] == lead ) { cnt[i]++; continue; }
			}
			else
			{
				/* with brush entities the stack is ordered by (key, then as the 1/z tests placed the pieces) */
				for( i = 0; i < depth; i++ ) if( stk[i] == lead ) break;
				if( i < depth ) { cnt[i]++; continue; }
				const int32_t ks = rc_span_key( surfs, lead );
				const int insub = ks != lead || ( surfs[lead - 1].emitted >> 8 ) != 0;
				i = 0;
				if( depth > 0 )
				{
					const int32_t k2 = rc_span_key( surfs, stk[0] );
					if( ks < k2 || ( insub && ks == k2 && rc_span_zi_in_front( &surfs[lead - 1], &surfs[stk[0] - 1], u, y )))
						i = 0;
					else
					{
						for( ;; )
						{
							do { i++; } while( i < depth && ks > rc_span_key( surfs, stk[i] ));
							if( i < depth && ks == rc_span_key( surfs, stk[i] ))
							{
								if( !insub ) continue;
								if( rc_span_zi_in_front( &surfs[lead - 1], &surfs[stk[i] - 1], u, y )) break;
								continue;
							}
							break;
						}
					}
				}
			}
			if( depth >= stkcap ) { *overflow = 1; continue; }
			if( i == 0 )
			{
				/* newtop: emit a span for the surface being obscured */
				RC_EMIT_SPAN( top, top_last_u, iu )
				top_last_u = iu;
			}
			for( int32_t k = depth; k > i; k-- ) { stk[k] = stk[k - 1]; cnt[k] = cnt[k - 1]; }
			stk[i] = lead; cnt[i] = 1; depth++;
			top = stk[0];
		}
	}
	/* R_CleanupSpan: whatever is on top runs to the right edge */
	{
		int32_t iu = tail_u >> 20;
		RC_EMIT_SPAN( top, top_last_u, iu )
	}
#undef RC_EMIT_SPAN
	return nspans;
}

/*
 * The same replay with the surface stack kept as a bit set.  Only two things about the key-ordered stack
 * ever reach the output: which surfaces are on it (spanstate >= 1) and which of them has the smallest key
 * (world keys are unique and equal the candidate order, so "smallest key" = lowest set bit).  A surface's
 * spanstate is 0 or 1 except on degenerate scanlines (inverted spans, a surface entered twice); those few
 * live in the exception list exc[] = {surf, spanstate} with spanstate outside {0, 1}.
 * keys: the line's events as sorted rc_event_key values.  bits: (maxsurf + 32) / 32 words, zeroed by the caller.
 * Returns the number of spans written.
 */
RC_HD int32_t rc_generate_spans_bits( const unsigned long long *keys, int32_t nev, int32_t head_iu, int32_t tail_u,
	uint32_t *bits, int32_t nwords, int32_t *exc, int32_t exccap, rc_span_t *spans, int32_t maxspans, int32_t *overflow )
{
	int32_t nexc = 0, nspans = 0;
	int32_t top = 0, top_last_u = head_iu; /* 0 = background (surfaces[1], key 0x7FFFFFFF) */
	unsigned long long k0 = nev > 0 ? keys[0] : 0ull, k1 = nev > 1 ? keys[1] : 0ull, k2 = nev > 2 ? keys[2] : 0ull, k3 = nev > 3 ? keys[3] : 0ull;

#define RC_EMIT_SPAN( SURF, LAST, IU ) \
	if(( IU ) > ( LAST )) { if( nspans < maxspans ) { spans[nspans].u = (int16_t)( LAST ); spans[nspans].count = (int16_t)(( IU ) - ( LAST )); spans[nspans].surf = ( SURF ); nspans++; } else *overflow = 1; }

	for( int32_t e = 0; e < nev; e++ )
	{
		/* keys are rc_event_key values: (u ^ sign bit) << 32 | surfaces; spans may overwrite the key array
		 * (span k is written after key k was read) */
		/* four keys in flight while earlier events are replayed (spans written so far sit at slots <= e) */
		const unsigned long long key = k0;
		k0 = k1; k1 = k2; k2 = k3;
		if( e + 4 < nev ) k3 = keys[e + 4];
		const int32_t u = (int32_t)((uint32_t)( key >> 32 ) ^ 0x80000000u );
		const uint32_t both = (uint32_t)key;
		if( u > tail_u )
			break;
		const int32_t iu = u >> 20;
		const int32_t trail = (int32_t)( both >> 16 ), lead = (int32_t)( both & 0xFFFF );

		if( trail )
		{
			int32_t x = -1;
			for( int32_t i = 0; i < nexc; i++ ) if( exc[2 * i] == trail ) { x = i; break; }
			if( x >= 0 )
			{
				/* spanstate 2 -> 1 returns to the plain representation (bit stays set); others just count down */
				if( --exc[2 * x + 1] == 1 ) { exc[2 * x] = exc[2 * ( nexc - 1 )]; exc[2 * x + 1] = exc[2 * ( nexc - 1 ) + 1]; nexc--; }
			}
			else if(( bits[trail >> 5] >> ( trail & 31 )) & 1 )
			{
				/* R_TrailingEdge: --spanstate == 0 -> leave the stack */
				bits[trail >> 5] &= ~( 1u << ( trail & 31 ));
				if( trail == top )
				{
					RC_EMIT_SPAN( trail, top_last_u, iu )
					top_last_u = iu;
					/* the next surface on the stack: lowest set bit above the old top */
					int32_t wi = trail >> 5;
					uint32_t wv = bits[wi] & ~(( 2u << ( trail & 31 )) - 1u );
					top = 0;
					for( ;; )
					{
						if( wv )
						{
#if defined( __CUDA_ARCH__ )
							top = ( wi << 5 ) + ( __ffs( (int)wv ) - 1 );
#else
							top = ( wi << 5 ) + __builtin_ctz( wv );
#endif
							break;
						}
						if( ++wi >= nwords ) break;
						wv = bits[wi];
					}
				}
			}
			else
			{
				/* inverted span: spanstate goes negative */
				if( nexc < exccap ) { exc[2 * nexc] = trail; exc[2 * nexc + 1] = -1; nexc++; }
				else *overflow = 1;
			}
		}
		if( lead )
		{
			int32_t x = -1;
			for( int32_t i = 0; i < nexc; i++ ) if( exc[2 * i] == lead ) { x = i; break; }
			if( x >= 0 )
			{
				/* spanstate -1 -> 0 returns to the plain representation (bit stays clear) */
				if( ++exc[2 * x + 1] == 0 ) { exc[2 * x] = exc[2 * ( nexc - 1 )]; exc[2 * x + 1] = exc[2 * ( nexc - 1 ) + 1]; nexc--; }
			}
			else if(( bits[lead >> 5] >> ( lead & 31 )) & 1 )
			{
				/* already on the stack: spanstate 2 */
				if( nexc < exccap ) { exc[2 * nexc] = lead; exc[2 * nexc + 1] = 2; nexc++; }
				else *overflow = 1;
			}
			else
			{
				/* R_LeadingEdge: ++spanstate == 1 -> enter the stack; in front of everything -> new top */
				bits[lead >> 5] |= 1u << ( lead & 31 );
				if( !top || lead < top )
				{
					RC_EMIT_SPAN( top, top_last_u, iu )
					top_last_u = iu;
					top = lead;
				}
			}
		}
	}
	{
		int32_t iu = tail_u >> 20;
		RC_EMIT_SPAN( top, top_last_u, iu )
	}
#undef RC_EMIT_SPAN
	return nspans;
}

/* ------------------------------------------------------------------------------------------ */
/* A12: translucent brush entities - R_DrawBrushModel (ref/soft/r_main.c:861): an edge frame of the entity's own   */
/* (R_BeginEdgeFrame .. R_ScanEdges with alphaspans), its pieces keyed by the leaves the world pass numbered        */

/* the entity's pieces in one view: surface records (mip 0, leaf keys) and edges; returns the piece count */
RC_HD int32_t rc_tbrush_pieces( const rc_scene_t *sc, const rc_view_t *vw, const rc_bent_t *bents, int32_t nbents, int32_t ent,
	const uint32_t *nodevis, const uint32_t *leafvis, const int32_t *leafbase, const int32_t *leafkey, const uint16_t *candof,
	rc_surf_t *surfs, rc_edge_t *edges, int32_t *nedges, uint32_t *cache_need, int32_t *overflow )
{
	const rc_bent_t *be = &bents[ent];
	*nedges = 0;
	const int32_t clipflags = rc_bmodel_check_bbox( vw, be->minmaxs );
	if( clipflags < 0 )
		return 0;
	const int32_t topnode = rc_find_topnode( sc, nodevis, leafvis, be->minmaxs );
	if( topnode == RC_TOPNODE_NONE )
		return 0;
	rc_view_t bv;
	rc_bmodel_view( vw, be, &bv );
	rc_bmout_t out;
	out.surfs = surfs; out.max_cands = RC_TB_MAXSURFS;
	out.edges = edges; out.max_edges = RC_TB_MAXEDGES; out.nedges = nedges;
	out.leafbase = leafbase; out.leafkey = leafkey; out.candof = candof; out.cache_need = cache_need;
	out.nsurfs = 0; out.overflow = 0; out.noutmax = 0;
	for( int32_t f = 0; f < be->numfaces; f++ )
		rc_bmodel_face( sc, &bv, bents, nbents, ent, be->firstface + f, clipflags, topnode, nodevis, leafvis, &out );
	if( out.overflow || out.noutmax > RC_TB_MAXEDGES ) *overflow = 1;
	if( *nedges > RC_TB_MAXEDGES ) *nedges = RC_TB_MAXEDGES;
	return out.nsurfs;
}

/* R_ScanEdges for one scanline of that edge frame: the events of the line, sorted, through the span generator */
#define RC_TB_LINEEVENTS 64
RC_HD int32_t rc_tbrush_line( const rc_surf_t *surfs, const rc_edge_t *edges, int32_t nedges, int32_t y, int32_t head_iu, int32_t tail_u,
	rc_span_t *spans, int32_t maxspans, int32_t *overflow )
{
	unsigned long long keys[RC_TB_LINEEVENTS];
	int32_t n = 0;
	for( int32_t i = 0; i < nedges; i++ )
	{
		if( !rc_edge_active( &edges[i], y ))
			continue;
		if( n >= RC_TB_LINEEVENTS ) { *overflow = 1; break; }
		const unsigned long long key = rc_event_key( rc_edge_u_at( &edges[i], y ), edges[i].surfs );
		int32_t j = n - 1;
		while( j >= 0 && keys[j] > key ) { keys[j + 1] = keys[j]; j--; }
		keys[j + 1] = key;
		n++;
	}
	if( !n )
		return 0;
	int32_t stk[RC_TB_MAXSURFS], cnt[RC_TB_MAXSURFS], neg[16];
	return rc_generate_spans( keys, n, head_iu, tail_u, (const uint16_t *)0, stk, cnt, RC_TB_MAXSURFS, neg, 8, spans, maxspans, overflow, surfs, y );
}

/* ------------------------------------------------------------------------------------------ */
/* A9/A10/A11: span drawing split into chunk set-up (float, once per 8 or 16 pixels) and the    */
/* per-pixel integer part (ref/soft/r_scan.c:89-221,528-668,1309-1357)                          */

enum { RC_KIND_BACKGROUND = 0, RC_KIND_SOLID = 1, RC_KIND_TURB = 2, RC_KIND_SKY = 3 };

typedef struct
{
	int32_t s, t, sstep, tstep; /* 16.16 texture coordinates at the chunk's first pixel + per-pixel step */
	int32_t izi, izistep;       /* 1/z (<<31) at the chunk's first pixel + per-pixel step */
	int32_t base, width;        /* texel offset and row pitch of the cache block / texture; sky: base = colour */
	int32_t u0, kind;
} rc_chunk_t; /* 40 B */

RC_HD int32_t rc_span_kind( const rc_surf_t *sf )
{
	/* a solid face without a cache block (lightmap sample size other than 16: R_NewMap reports them) is filled flat
	 * like a hole in the map - the cache arena is never indexed with its "no block" offset */
	if(( sf->flags & RC_SURF_DRAWSKY ) || sf->cache_ofs < 0 ) return RC_KIND_SKY;
	if( sf->flags & RC_SURF_DRAWTURB ) return RC_KIND_TURB;
	return RC_KIND_SOLID;
}

RC_HD int32_t rc_span_nchunks( int32_t kind, int32_t count )
{
	return kind == RC_KIND_TURB ? ( count + 15 ) >> 4 : ( count + 7 ) >> 3;
}

RC_HD int32_t rc_clamp_st( int32_t x, int32_t hi, int32_t lo )
{
	if( x > hi ) x = hi;
	else if( x < lo ) x = lo;
	return x;
}

/* D_DrawZSpans (r_scan.c:1309): 1/z of the span's first pixel as a 1.31 integer + the per-pixel step; later
 * pixels are integer steps from there (int32 wraparound) */
RC_HD void rc_span_z0( const rc_surf_t *sf, int32_t u, int32_t v, int32_t *izi0, int32_t *izistep )
{
	float du = (float)u, dv = (float)v;
	float zi = sf->d_ziorigin + dv * sf->d_zistepv + du * sf->d_zistepu;
	*izistep = rc_f2i( sf->d_zistepu * (float)0x8000 * (float)0x10000 );
	*izi0 = rc_f2i( zi * (float)0x8000 * (float)0x10000 );
}

/* texture coordinates (16.16) of the first pixel of chunk j of the span (u, v, count) and their per-pixel step */
RC_HD void rc_chunk_st( const rc_surf_t *sf, int32_t kind, int32_t u, int32_t v, int32_t count, int32_t j,
	int32_t *ps, int32_t *pt, int32_t *psstep, int32_t *ptstep )
{
	int32_t px = kind == RC_KIND_TURB ? 16 : 8;
	float du = (float)u, dv = (float)v;
	float zi = sf->d_ziorigin + dv * sf->d_zistepv + du * sf->d_zistepu;
	float sdivz = sf->d_sdivzorigin + dv * sf->d_sdivzstepv + du * sf->d_sdivzstepu;
	float tdivz = sf->d_tdivzorigin + dv * sf->d_tdivzstepv + du * sf->d_tdivzstepu;
	float fpx = (float)px;
	float sdivzstep = sf->d_sdivzstepu * fpx, tdivzstep = sf->d_tdivzstepu * fpx, zistep = sf->d_zistepu * fpx;
	int32_t nchunks = rc_span_nchunks( kind, count );
	int32_t sstep = 0, tstep = 0;

	/* s/t of the span's first pixel: D_DrawSpans16 never reloads s from snext, it keeps stepping
	 * (s += sstep eight times, sstep = (snext - s) >> 3), so every later chunk start is congruent to
	 * s0 modulo 8: s_j = snext_(j-1) - ((snext_(j-1) - s0) & 7).  Turbulent8 does reload (r_turb_s = snext). */
	float z = (float)0x10000 / zi;
	int32_t s0 = rc_clamp_st( rc_f2i( sdivz * z ) + sf->sadjust, sf->bbextents, 0 );
	int32_t t0 = rc_clamp_st( rc_f2i( tdivz * z ) + sf->tadjust, sf->bbextentt, 0 );
	int32_t s = s0, t = t0;
	if( j )
	{
		for( int32_t i = 0; i < j; i++ )
		{
			sdivz += sdivzstep; tdivz += tdivzstep; zi += zistep;
		}
		z = (float)0x10000 / zi;
		s = rc_clamp_st( rc_f2i( sdivz * z ) + sf->sadjust, sf->bbextents, px );
		t = rc_clamp_st( rc_f2i( tdivz * z ) + sf->tadjust, sf->bbextentt, px );
		if( kind != RC_KIND_TURB )
		{
			s -= ( s - s0 ) & 7;
			t -= ( t - t0 ) & 7;
		}
	}

	int32_t spancount = count - j * px;
	if( j < nchunks - 1 )
	{
		sdivz += sdivzstep; tdivz += tdivzstep; zi += zistep;
		z = (float)0x10000 / zi;
		int32_t snext = rc_clamp_st( rc_f2i( sdivz * z ) + sf->sadjust, sf->bbextents, px );
		int32_t tnext = rc_clamp_st( rc_f2i( tdivz * z ) + sf->tadjust, sf->bbextentt, px );
		int sh = kind == RC_KIND_TURB ? 4 : 3;
		sstep = ( snext - s ) >> sh;
		tstep = ( tnext - t ) >> sh;
	}
	else if( spancount > 1 )
	{
		float spancountminus1 = (float)( spancount - 1 );
		sdivz += sf->d_sdivzstepu * spancountminus1;
		tdivz += sf->d_tdivzstepu * spancountminus1;
		zi += sf->d_zistepu * spancountminus1;
		z = (float)0x10000 / zi;
		int32_t snext = rc_clamp_st( rc_f2i( sdivz * z ) + sf->sadjust, sf->bbextents, px );
		int32_t tnext = rc_clamp_st( rc_f2i( tdivz * z ) + sf->tadjust, sf->bbextentt, px );
		sstep = ( snext - s ) / ( spancount - 1 );
		tstep = ( tnext - t ) / ( spancount - 1 );
	}
	if( kind == RC_KIND_TURB )
	{
		s &= ( RC_CYCLE << 16 ) - 1;
		t &= ( RC_CYCLE << 16 ) - 1;
	}
	*ps = s; *pt = t; *psstep = sstep; *ptstep = tstep;
}

/*
 * State of D_DrawSpans16 / Turbulent8 at the start of chunk j of a span (u, v, count).  The reference
 * walks the chunks sequentially accumulating sdivz/tdivz/zi in float; chunk j repeats those j additions
 * so that chunks can be set up independently (one lane each) with bit-identical results.
 */
RC_HD void rc_chunk_setup( const rc_view_t *vw, const rc_surf_t *sf, int32_t sky_color, int32_t u, int32_t v, int32_t count, int32_t j, rc_chunk_t *c )
{
	int32_t kind = rc_span_kind( sf );
	int32_t px = kind == RC_KIND_TURB ? 16 : 8;
	int32_t izi0, izistep;
	(void)vw;
	rc_span_z0( sf, u, v, &izi0, &izistep );
	c->izistep = izistep;
	c->izi = (int32_t)((uint32_t)izi0 + (uint32_t)( j * px ) * (uint32_t)izistep );
	c->u0 = u + j * px;
	c->kind = kind;
	c->base = sf->cache_ofs;
	c->width = sf->cachewidth;
	c->s = c->t = c->sstep = c->tstep = 0;
	if( kind == RC_KIND_SKY )
	{
		c->base = sky_color;
		return;
	}
	rc_chunk_st( sf, kind, u, v, count, j, &c->s, &c->t, &c->sstep, &c->tstep );
}

/* texel index of pixel i (0-based inside the chunk) - "span texel indices must be integer-exact" */
RC_HD int32_t rc_pixel_texel( const rc_chunk_t *c, int32_t i, const int32_t *turb )
{
	int32_t s = (int32_t)((uint32_t)c->s + (uint32_t)i * (uint32_t)c->sstep );
	int32_t t = (int32_t)((uint32_t)c->t + (uint32_t)i * (uint32_t)c->tstep );
	if( c->kind == RC_KIND_TURB )
	{
		int32_t sturb = (( s + turb[( t >> 16 ) & ( RC_CYCLE - 1 )] ) >> 16 ) & 63;
		int32_t tturb = (( t + turb[( s >> 16 ) & ( RC_CYCLE - 1 )] ) >> 16 ) & 63;
		return c->base + ( tturb << 6 ) + sturb;
	}
	return c->base + ( s >> 16 ) + ( t >> 16 ) * c->width;
}

/* sw_texfilt 1 (D_DrawSpans16's second drawing loop, r_scan.c:652-672): the texel coordinates of a solid pixel are
 * nudged by a 2x2 ordered-dither kernel before they are truncated.  X = ( pspan->u + spancount ) & 1 with spancount the
 * pixels left in the 8-pixel chunk, pspan->u the SPAN's first column (same parity as the chunk's); Y = pspan->v & 1. */
RC_HD int32_t rc_texfilt_index( int32_t s, int32_t t, int32_t width, int32_t X, int32_t Y )
{
	/* kernel[X][Y][0 / 1] = { { {16384, 0}, {49152, 32768} }, { {32768, 49152}, {0, 16384} } } */
	const int32_t q = ( X << 1 ) | Y;
	const int32_t ks = q == 0 ? 16384 : q == 1 ? 49152 : q == 2 ? 32768 : 0;
	const int32_t kt = q == 0 ? 0 : q == 1 ? 32768 : q == 2 ? 49152 : 16384;
	int32_t is = (int32_t)((uint32_t)s + (uint32_t)ks ) >> 16, it = (int32_t)((uint32_t)t + (uint32_t)kt ) >> 16;
	is = is ? is - 1 : is;
	it = it ? it - 1 : it;
	return is + it * width;
}

/* pixel i of an n-pixel chunk of a solid span that starts at column u_span of scanline v */
RC_HD int32_t rc_pixel_texel_filt( const rc_chunk_t *c, int32_t i, int32_t n, int32_t u_span, int32_t v )
{
	int32_t s = (int32_t)((uint32_t)c->s + (uint32_t)i * (uint32_t)c->sstep );
	int32_t t = (int32_t)((uint32_t)c->t + (uint32_t)i * (uint32_t)c->tstep );
	return c->base + rc_texfilt_index( s, t, c->width, ( u_span + ( n - i )) & 1, v & 1 );
}

RC_HD int16_t rc_pixel_z( const rc_chunk_t *c, int32_t i )
{
	return (int16_t)((int32_t)((uint32_t)c->izi + (uint32_t)i * (uint32_t)c->izistep ) >> 16 );
}

/* ------------------------------------------------------------------------------------------ */
/* A7/A8: light map blend and surface cache block (ref/soft/r_surf.c:174-229,291-757)            */

/*
 * R_AddDynamicLights (ref/soft/r_surf.c:72-165) for luxel (s, t) of a world face: what the lights whose bit is set
 * in `bits` add to blocklights[s + t * smax].  World faces only (tr.modelviewIdentity), sample_frac 1 (no
 * TEX_WORLD_LUXELS).  The per-light constants are recomputed per luxel; float expressions in the reference's order.
 */
RC_HD uint32_t rc_dlight_luxel( const rc_scene_t *sc, const rc_surface_t *sf, const rc_dlight_t *lights, int32_t nlights, uint32_t bits, int32_t s, int32_t t )
{
	const rc_plane_t *pl = &sc->planes[sf->plane];
	const float sample_size = (float)sf->sample_size;
	uint32_t add = 0;
	for( int32_t k = 0; k < nlights; k++ )
	{
		const rc_dlight_t *dl = &lights[k];
		if( !( bits & ( 1u << dl->bit )))
			continue;
		float rad = dl->radius;
		const float dist = rc_planediff( dl->origin, pl );
		rad -= dist < 0 ? -dist : dist;
		float minlight = dl->minlight;
		if( rad < minlight )
			continue;
		minlight = rad - minlight;
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
		const float sl = rc_dot3( impact, sf->lmvecs[0] ) + sf->lmvecs[0][3] - (float)sf->lightmapmins[0];
		const float tl = rc_dot3( impact, sf->lmvecs[1] ) + sf->lmvecs[1][3] - (float)sf->lightmapmins[1];
		int32_t td = rc_f2i(( tl - sample_size * (float)t ) * 1.0f );
		if( td < 0 ) td = -td;
		int32_t sd = rc_f2i(( sl - sample_size * (float)s ) * 1.0f );
		if( sd < 0 ) sd = -sd;
		const float d = sd > td ? (float)( sd + ( td >> 1 )) : (float)( td + ( sd >> 1 ));
		if( d < minlight )
			add += (uint32_t)(( rc_f2i(( rad - d ) * 256.0f ) * dl->monolight ) / 256 );
	}
	return add;
}

/* R_BuildLightMap for luxel i of a face: sum of (r+g+b) * style value, dynamic lights (`extra`, rc_dlight_luxel),
 * gamma, quantise to bits 8..12 */
RC_HD uint32_t rc_blocklight( const rc_scene_t *sc, const rc_surface_t *sf, const int32_t *stylevalue, int32_t size, int32_t i, uint32_t extra )
{
	uint32_t bl = 0;
	if( sf->lightofs >= 0 )
	{
		for( int map = 0; map < RC_MAXLIGHTMAPS; map++ )
		{
			if( sf->styles[map] >= 255 )
				break;
			uint32_t scale = (uint32_t)stylevalue[sf->styles[map]];
			const uint8_t *lm = sc->lightdata + 3 * ((size_t)sf->lightofs + (size_t)map * size + i );
			bl += ((uint32_t)lm[0] + lm[1] + lm[2] ) * scale;
		}
	}
	bl += extra;
	int32_t t;
	if( bl < 65280 )
	{
		uint32_t b = bl >> 6;
		t = (int32_t)(( b >= 1024 ? 0u : (uint32_t)sc->lightgamma[b] ) << 6 );
	}
	else t = (int32_t)bl;
	t = t >= 0 ? ( t < 65535 * 3 ? t : 65535 * 3 ) : 0; /* bound( 0, t, 65535 * 3 ) */
	t = t / 2048 / 3;
	return (uint32_t)( t << 8 );
}

/*
 * One texel (x, y) of the cache block of (face, mip): R_DrawSurface + R_DrawSurfaceBlock8_mipN in
 * closed form.  blocklights = the face's luxel grid (lightwidth per row) from rc_blocklight.
 * All light arithmetic is uint32 with the reference's wraparound and unsigned shifts.
 */
RC_HD uint16_t rc_cache_texel( const rc_scene_t *sc, const rc_surface_t *sf, const rc_texture_t *tex, int32_t mip,
	const uint32_t *blocklights, int32_t lightwidth, int32_t x, int32_t y )
{
	int32_t k = 4 - mip, bs = 16 >> mip;
	int32_t bu = x >> k, b = x & ( bs - 1 ), bv = y >> k, i = y & ( bs - 1 );

	const uint32_t *lp = blocklights + bv * lightwidth + bu;
	uint32_t ll = lp[0], lr = lp[1];
	uint32_t lls = ( lp[lightwidth] - ll ) >> k, lrs = ( lp[lightwidth + 1] - lr ) >> k;
	ll += (uint32_t)i * lls; lr += (uint32_t)i * lrs;
	uint32_t step = ( ll - lr ) >> k;
	uint32_t light = lr + (uint32_t)( bs - 1 - b ) * step;

	/* source texel: column offset advances per block column and snaps to 0 past the texture width,
	 * row offset advances per block row and steps back by the texture height (r_surf.c:411-426,586-600) */
	int32_t smax = tex->width >> mip, tmax = tex->height >> mip;
	if( smax < 1 ) smax = 1;
	if( tmax < 1 ) tmax = 1;
	int32_t soffset = ((( (int32_t)sf->lightmapmins[0] ) >> mip ) + ( smax << 16 )) % smax;
	for( int32_t q = 0; q < bu; q++ )
	{
		soffset += bs;
		if( soffset >= smax ) soffset = 0;
	}
	int32_t rowofs = (((( (int32_t)sf->lightmapmins[1] ) >> mip ) + ( tmax << 16 )) % tmax ) * smax;
	int32_t total = tmax * smax;
	for( int32_t q = 0; q < bv; q++ )
	{
		rowofs += bs * smax;
		if( rowofs + soffset >= total ) rowofs -= total; /* psource >= r_sourcemax */
	}
	int32_t idx = rowofs + i * smax + soffset + b;
	/* the reference reads past the mip when a texture is not a multiple of the block size; stay inside the arena */
	int32_t lim = tex->width * tex->height * 2;
	if( idx < 0 ) idx = 0;
	if( idx >= lim ) idx = lim - 1;
	uint16_t pix = sc->texels[tex->ofs[mip] + idx];
	uint16_t out = (uint16_t)( sc->colormap[( pix >> 3 ) | (( light & 0x1f00 ) << 5 )] | ( pix & 7 ));
	if( mip == 0 && pix == RC_TRANSPARENT_COLOR )
		out = RC_TRANSPARENT_COLOR;
	return out;
}

#endif /* RC_CORE_H */
