/*
 * rc_k_tris.cuh - S5, entities: everything ref_soft draws that is not a world face goes through its triangle
 * rasteriser (R_DrawEntitiesOnList -> TriAPI -> R_RenderTriangle -> R_DrawTriangle, ref/soft/r_main.c:504,
 * r_triapi.c:221, r_trialias.c:171, r_polyse.c:636), after the world's spans are in the frame.  The arithmetic is
 * rc_tri.h; here is the decomposition:
 *
 *   k_item_order  block per view: the order the view's draw items are drawn in - the fixed ones as the host listed them,
 *                 the translucent entities sorted far to near by R_TransEntityCompare (r_main.c:157) for THIS eye.
 *   k_tri_polys   thread per (view, triangle): world-space corners (recorded TriAPI corners, or a sprite's quad built
 *                 from the view's axes, r_sprite.c:120-266) -> R_SetupFinalVert, trivial reject, R_AliasClipTriangle ->
 *                 the screen-space polygon the reference would fan into triangles, left in HBM.
 *   k_tris_ordered  the span drawers read and write the frame, and all but the opaque one blend: triangles have to
 *                 reach a pixel in submission order.  Block per (view, horizontal strip of the frame): it walks the
 *                 view's items in order, 128 at a time; between two translucent brush entities the items' polygons are
 *                 one flat list.  Thread = polygon: the fan triangles whose scanlines reach into the strip, compacted
 *                 in order (block prefix sum); the others cost nothing more.  Those are set up
 *                 TR_BATCH at a time (thread = triangle, rc_tri_setup), then drawn one after the other: thread =
 *                 scanline (closed form, rc_tri_row), a prefix sum over the scanlines' pixel counts, thread = pixel.
 *                 Every scanline and pixel is closed form, so the result is the reference's sequential one.
 */
#ifndef RC_K_TRIS_CUH
#define RC_K_TRIS_CUH

#include "rc_tri.h"

#define TR_THREADS 128
#define TR_STRIPS  4     /* horizontal strips of a frame drawn by different blocks */

struct rc_poly_t
{
	int32_t k;           /* vertices, 0 = nothing to draw */
	int32_t v[8][6];     /* u, v, s, t, l, zi */
}; /* 196 B */

/* the order the items of one view are drawn in */
__global__ void __launch_bounds__( 128 ) k_item_order( rc_work_t w )
{
	const int view = blockIdx.x, t = threadIdx.x;
	const rc_view_t *vw = &w.views[view];
	int32_t *order = w.item_order + (size_t)view * w.nitems;
	__shared__ int32_t s_pos[2048]; /* MAX_VISIBLE_PACKET translucent entities: position of entity e in the sorted list */
	const int32_t ns = w.nsortents < 2048 ? w.nsortents : 2048;
	for( int32_t a = t; a < ns; a += blockDim.x )
	{
		const rc_sortent_t sa = w.sortents[a];
		float v[3] = { vw->origin[0] - sa.center[0], vw->origin[1] - sa.center[1], vw->origin[2] - sa.center[2] };
		const float da = sa.use_dist ? rc_dot3( v, v ) : 1000000000.0f;
		int32_t before = 0;
		for( int32_t b = 0; b < ns; b++ )
		{
			const rc_sortent_t sb = w.sortents[b];
			float u[3] = { vw->origin[0] - sb.center[0], vw->origin[1] - sb.center[1], vw->origin[2] - sb.center[2] };
			const float db = sb.use_dist ? rc_dot3( u, u ) : 1000000000.0f;
			/* farther first, then the lower render-mode rank; equal keys keep the list order (merge sort is stable) */
			if( db > da || ( db == da && ( sb.rank < sa.rank || ( sb.rank == sa.rank && b < a )))) before++;
		}
		s_pos[before] = a;
	}
	__syncthreads();
	if( t == 0 )
	{
		int32_t n = 0;
		bool trans_done = false;
		for( int32_t i = 0; i < w.nitems; i++ )
		{
			if( w.items[i].sortent < 0 ) { order[n++] = i; continue; }
			if( trans_done ) continue;
			trans_done = true;
			for( int32_t p = 0; p < ns; p++ )
			{
				const rc_sortent_t se = w.sortents[s_pos[p]];
				for( int32_t j = 0; j < se.nitems; j++ ) order[n++] = se.first_item + j;
			}
		}
		for( ; n < w.nitems; n++ ) order[n] = -1;
	}
}

/* world-space corners of triangle tri of an item for a view; 0 = not drawn */
__device__ __forceinline__ int rc_item_corners( const rc_work_t &w, const rc_view_t *vw, const rc_item_t *it, int32_t tri, rc_corner_t *c )
{
	if( rc_item_is_quad( it ))
	{
		rc_corner_t q[4];
		rc_item_t gi;
		if( it->glow >= 0 )
		{
			/* a glow sprite: this view's verdict, size and brightness */
			const rc_glow_t g = w.glow[(size_t)( vw - w.views ) * w.nglow + it->glow];
			if( !g.visible )
				return 0;
			gi = *it; gi.scale = g.scale; gi.light = g.light;
			it = &gi;
		}
		if( !rc_quad_corners( vw, it, q ))
			return 0;
		c[0] = q[0];
		c[1] = tri ? q[2] : q[1];
		c[2] = tri ? q[3] : q[2];
		return 1;
	}
	if( it->kind == RC_ITEM_MESH )
	{
		/* a studio mesh: static strip corners, the entity's transformed vertices and per-normal lights (k_studio_points) */
		const rc_mcorner_t *mc = reinterpret_cast<const rc_mcorner_t *>( w.sstatic ) + it->first + 3 * tri;
#pragma unroll
		for( int q = 0; q < 3; q++ )
		{
			/* one 8-byte load for the corner record, one 16-byte load for the vertex (k_studio_points leaves float4s) */
			union { uint2 raw; rc_mcorner_t m; } u;
			u.raw = __ldg( reinterpret_cast<const uint2 *>( mc + q ));
			const rc_mcorner_t m = u.m;
			const float4 p = __ldg( reinterpret_cast<const float4 *>( w.mvpos ) + ( it->vert_base + m.vert ));
			c[q].xyz[0] = p.x; c[q].xyz[1] = p.y; c[q].xyz[2] = p.z;
			c[q].l = w.mlight[it->light_base + m.norm]; c[q].s = m.s; c[q].t = m.t;
			if( it->chrome )
			{
				const rc_spart_t *pt = &w.parts[it->part];
				int32_t cs, ct;
				rc_studio_chrome_st( vw, w.bones + (size_t)( pt->bones + w.sstatic[pt->snormbone + m.norm] ) * 12,
					reinterpret_cast<const float *>( w.sstatic ) + pt->snorms + 3 * m.norm, it->skinwidth, it->skinheight, &cs, &ct );
				c[q].s = (int16_t)cs; c[q].t = (int16_t)ct;
			}
		}
		return 1;
	}
	const rc_corner_t *src = w.corners + it->first + 3 * tri;
	c[0] = src[0]; c[1] = src[1]; c[2] = src[2];
	return 1;
}

/* the item a triangle number belongs to: items are few, their order0 ascending */
__device__ __forceinline__ const rc_item_t *rc_item_of( const rc_work_t &w, int32_t g )
{
	int32_t lo = 0, hi = w.nitems - 1;
	while( lo < hi )
	{
		const int32_t mid = ( lo + hi + 1 ) >> 1;
		if( w.items[mid].order0 <= g ) lo = mid; else hi = mid - 1;
	}
	return &w.items[lo];
}

__global__ void k_partbox_init( rc_work_t w, int32_t nparts )
{
	const int32_t g = blockIdx.x * blockDim.x + threadIdx.x;
	if( g < nparts * 6 ) w.partbox[g] = ( g % 6 ) < 3 ? 0x7FFFFFFF : (int32_t)0x80000000;
}

/* R_StudioDrawPoints' vertex and normal loops (r_studio.c:1846-1895) for every drawn (entity, submodel) of the batch, once
 * per batch (they do not depend on the view): thread per vertex, then thread per normal */
__global__ void __launch_bounds__( 256 ) k_studio_points( rc_scene_t sc, rc_work_t w, int32_t nparts, int32_t totalverts, int32_t totalnorms )
{
	const int32_t g = blockIdx.x * blockDim.x + threadIdx.x;
	if( g >= totalverts + totalnorms )
		return;
	const bool isnorm = g >= totalverts;
	const int32_t idx = isnorm ? g - totalverts : g;
	/* the part whose slice holds idx */
	int32_t lo = 0, hi = nparts - 1;
	while( lo < hi )
	{
		const int32_t mid = ( lo + hi + 1 ) >> 1;
		if(( isnorm ? w.parts[mid].light_base : w.parts[mid].vert_base ) <= idx ) lo = mid; else hi = mid - 1;
	}
	const rc_spart_t pt = w.parts[lo];
	const float *sf = reinterpret_cast<const float *>( w.sstatic );
	if( !isnorm )
	{
		const int32_t i = idx - pt.vert_base;
		if( i >= pt.numverts ) return;
		float out[3];
		rc_studio_vertex( w.bones + (size_t)( pt.bones + w.sstatic[pt.svertbone + i] ) * 12, sf + pt.sverts + 3 * i, out );
		reinterpret_cast<float4 *>( w.mvpos )[idx] = make_float4( out[0], out[1], out[2], 0.0f );
		/* the part's bounding box for k_item_cull: order-preserving int keys, one atomic pair per axis (per warp when the
		 * whole warp works on one part) */
		const unsigned am = __activemask();
		const bool uniform = __match_any_sync( am, lo ) == am;
#pragma unroll
		for( int a = 0; a < 3; a++ )
		{
			const int32_t b = __float_as_int( out[a] ), key = b >= 0 ? b : b ^ 0x7FFFFFFF;
			int32_t kmin = key, kmax = key;
			if( uniform ) { kmin = __reduce_min_sync( am, key ); kmax = __reduce_max_sync( am, key ); }
			if( !uniform || ( threadIdx.x & 31 ) == __ffs( (int)am ) - 1 )
			{
				atomicMin( &w.partbox[lo * 6 + a], kmin );
				atomicMax( &w.partbox[lo * 6 + 3 + a], kmax );
			}
		}
	}
	else
	{
		const int32_t k = idx - pt.light_base;
		if( k >= pt.numnorms ) return;
		w.mlight[idx] = rc_studio_light( &pt, w.sstatic[pt.snormflags + k], sf + pt.snorms + 3 * k,
			w.blight + (size_t)( pt.blightvec + w.sstatic[pt.snormbone + k] ) * 3, sc.lightgamma );
	}
}

/*
 * Per (view, item): can any triangle of the item be drawn?  A triangle is dropped when its three final verts share a
 * clip flag (R_RenderTriangle's trivial reject, rc_view_triangle); every vertex of a studio part lies in the part's box,
 * and "z < ALIAS_Z_CLIP_PLANE", "u < vrect.x", ... are half-spaces of view space, so a box whose eight corners lie in
 * one of them (with a pixel of margin for the truncation of u, v) holds only vertices that carry that flag - also the
 * ones the clipper would make on its edges.  R_StudioDrawModel's own R_StudioCheckBBox plays this part in the reference.
 */
__global__ void __launch_bounds__( 128 ) k_item_cull( rc_work_t w )
{
	const int32_t g = blockIdx.x * blockDim.x + threadIdx.x;
	if( g >= w.nviews * w.nitems )
		return;
	const int view = g / w.nitems, idx = g - view * w.nitems;
	const rc_item_t *it = &w.items[idx];
	uint8_t vis = 1;
	if( it->kind == RC_ITEM_MESH && it->part >= 0 )
	{
		const rc_view_t *vw = &w.views[view];
		float bb[6];
#pragma unroll
		for( int a = 0; a < 6; a++ )
		{
			const int32_t key = w.partbox[it->part * 6 + a];
			bb[a] = __int_as_float( key >= 0 ? key : key ^ 0x7FFFFFFF );
		}
		const float xl = vw->xcenter - (float)vw->vrect[0] + 2.0f, xr = vw->xcenter - (float)( vw->vrect[0] + vw->vrect[2] ) - 2.0f;
		const float yt = vw->ycenter - (float)vw->vrect[1] + 2.0f, yb = vw->ycenter - (float)( vw->vrect[1] + vw->vrect[3] ) - 2.0f;
		unsigned all = 0x1Fu;
#pragma unroll
		for( int c = 0; c < 8; c++ )
		{
			const float p[3] = { bb[( c & 1 ) ? 3 : 0], bb[( c & 2 ) ? 4 : 1], bb[( c & 4 ) ? 5 : 2] };
			const float x = rc_dot3( p, vw->aliastransform[0] ) + vw->aliastransform[0][3];
			const float y = rc_dot3( p, vw->aliastransform[1] ) + vw->aliastransform[1][3];
			const float z = rc_dot3( p, vw->aliastransform[2] ) + vw->aliastransform[2][3];
			const float az = fabsf( z ) * 1e-4f + 1e-3f; /* slack for the rounding of the three products */
			unsigned f = 0;
			if( z < RC_ALIAS_Z_CLIP_PLANE - 0.25f ) f |= 1u;
			if( x * vw->xscale + xl * z < -az * vw->xscale ) f |= 2u;  /* u < vrect.x for every z > 0 */
			if( x * vw->xscale + xr * z > az * vw->xscale ) f |= 4u;   /* u > vrect.x + vrect.width */
			if( y * vw->yscale + yt * z < -az * vw->yscale ) f |= 8u;
			if( y * vw->yscale + yb * z > az * vw->yscale ) f |= 16u;
			all &= f;
		}
		if( all ) vis = 0;
	}
	w.itemvis[g] = vis;
	/* the opaque run: k_tris_fast is launched over the pairs that are left */
	if( vis && idx < w.nfast_items && it->tex_ofs >= 0 && it->count > 0 )
		w.pairs[atomicAdd( w.pair_count, 1 )] = ( view << 16 ) | idx;
}

__global__ void __launch_bounds__( 128 ) k_tri_polys( rc_work_t w, int32_t ntris )
{
	const int view = blockIdx.y;
	const int32_t gl = blockIdx.x * blockDim.x + threadIdx.x;
	if( gl >= ntris )
		return;
	const int32_t g = gl + w.fast_tris; /* the triangles before fast_tris go through k_tris_fast */
	const rc_item_t *it = rc_item_of( w, g );
	const rc_view_t *vw = &w.views[view];
	rc_poly_t *out = w.polys + (size_t)view * ntris + gl;
	rc_corner_t c[3];
	rc_fv_t poly[8];
	int k = 0;
	if( it->tex_ofs >= 0 && g - it->order0 < it->count && w.itemvis[(size_t)view * w.nitems + ( it - w.items )]
		&& rc_item_corners( w, vw, it, g - it->order0, c ))
		k = rc_view_triangle( vw, &c[0], &c[1], &c[2], poly );
	out->k = k;
	for( int i = 0; i < k; i++ )
	{
		out->v[i][0] = poly[i].u; out->v[i][1] = poly[i].v; out->v[i][2] = poly[i].s;
		out->v[i][3] = poly[i].t; out->v[i][4] = poly[i].l; out->v[i][5] = poly[i].zi;
	}
	/* RC_CTR_TRIS: the fan triangles R_DrawTriangle accepts (k_tris_ordered sets up only those that reach into a strip) */
	int32_t ndrawn = 0;
	rc_trisetup_t ts;
	for( int i = 1; i < k - 1; i++ )
		ndrawn += rc_tri_setup( &poly[0], &poly[i], &poly[i + 1], it->skinwidth, it->skinheight, &ts );
	if( ndrawn ) atomicAdd( &w.counters[RC_CTR_TRIS], (unsigned long long)ndrawn );
}

/*
 * A translucent brush entity in its turn (R_DrawBrushModel's R_ScanEdges + D_AlphaSurf): the warps of the block take the
 * strip's scanlines 32 at a time.  Lane = scanline: the line's events out of the entity's edges, sorted, through the span
 * generator (rc_tbrush_line) into shared memory; then line by line, lane = 8-pixel (16 for water) subdivision of a span:
 * the reference's float chain to the subdivision (rc_chunk_setup), its pixels through D_AlphaSpans16 / D_BlendSpans16 /
 * D_AddSpans16 / D_DrawTurbulent8ZSpan (rc_alpha_pixel).  A pixel belongs to one span of the frame: no two lanes meet.
 */
__device__ __forceinline__ void rc_tbrush_draw( const rc_scene_t &sc, const rc_work_t &w, const rc_blendluts_t *L, int view, int32_t ent,
	int32_t ys0, int32_t ys1, uint16_t *color, int16_t *zb, rc_span_t *spans, int32_t *nspans, int lane, int warp )
{
	const int32_t tb = w.tbord[ent];
	if( tb < 0 )
		return;
	const size_t slot = (size_t)view * w.ntb + tb;
	const int32_t ns = w.tb_counts[2 * slot], nedges = w.tb_counts[2 * slot + 1];
	if( ns <= 0 || nedges <= 0 )
		return;
	const rc_bent_t *be = &w.bents[ent];
	const rc_surf_t *surfs = w.tb_surfs + slot * RC_TB_MAXSURFS;
	const rc_edge_t *edges = w.tb_edges + slot * RC_TB_MAXEDGES;
	const rc_view_t *vw = &w.views[view];
	const int32_t *turb = sc.sintable + vw->turb_phase;
	const int32_t x0 = vw->vrect[0], vy0 = vw->vrect[1], vy1 = vw->vrect[1] + vw->vrect[3];
	const int32_t head_iu = x0, tail_u = (( x0 + vw->vrect[2] ) << 20 ) + 0xFFFFF;
	const int32_t ya = max( ys0, vy0 ), yb = min( ys1, vy1 );
	const uint32_t alpha = (uint32_t)min( be->alpha, 7 );
	for( int32_t yg = ya + warp * 32; yg < yb; yg += ( TR_THREADS / 32 ) * 32 )
	{
		const int32_t y = yg + lane;
		int32_t n = 0, ovf = 0;
		if( y < yb )
			n = rc_tbrush_line( surfs, edges, nedges, y, head_iu, tail_u, spans + lane * RC_TB_LINESPANS, RC_TB_LINESPANS, &ovf );
		nspans[lane] = n;
		if( ovf ) atomicAdd( &w.counters[RC_CTR_OVERFLOW_EVENTS], 1ull );
		__syncwarp();
		for( int32_t ln = 0; ln < 32 && yg + ln < yb; ln++ )
		{
			const int32_t yl = yg + ln;
			for( int32_t si = 0; si < nspans[ln]; si++ )
			{
				const rc_span_t sp = spans[ln * RC_TB_LINESPANS + si];
				if( !sp.surf )
					continue;
				const rc_surf_t *sf = &surfs[sp.surf - 1];
				const int32_t kind = rc_span_kind( sf );
				if( kind == RC_KIND_SKY || ( kind == RC_KIND_TURB && alpha == 0 ))
					continue;
				const int32_t nch = rc_span_nchunks( kind, sp.count ), cpx = kind == RC_KIND_TURB ? 16 : 8;
				for( int32_t j = lane; j < nch; j += 32 )
				{
					rc_chunk_t c;
					rc_chunk_setup( vw, sf, vw->sky_color, sp.u, yl, sp.count, j, &c );
					const int32_t npx = min( cpx, sp.count - j * cpx );
					const size_t base = (size_t)yl * w.width + c.u0;
					for( int32_t i = 0; i < npx; i++ )
					{
						const int32_t ti = rc_pixel_texel( &c, i, turb );
						const uint32_t texel = kind == RC_KIND_TURB ? __ldg( sc.texels + ti ) : sc.cache[ti];
						const int32_t izi = (int32_t)((uint32_t)c.izi + (uint32_t)i * (uint32_t)c.izistep );
						rc_alpha_pixel( L, be->trans, alpha, kind == RC_KIND_TURB, texel, izi, &color[base + i], &zb[base + i] );
					}
				}
			}
		}
		__syncwarp();
	}
}

#define TR_BATCH 16 /* triangles of a strip set up side by side before they are drawn one after the other */

/* block-wide exclusive prefix sum of one value per thread; *total = the sum.  Two barriers. */
__device__ __forceinline__ int32_t rc_block_excl_scan( int32_t v, int32_t *wsum, int t, int32_t *total )
{
	const int lane = t & 31, warp = t >> 5;
	const int32_t incl = rc_warp_incl_scan( v, lane );
	__syncthreads(); /* wsum of the previous scan has been read */
	if( lane == 31 ) wsum[warp] = incl;
	__syncthreads();
	int32_t wbase = 0, sum = 0;
#pragma unroll
	for( int k = 0; k < TR_THREADS / 32; k++ ) { if( k < warp ) wbase += wsum[k]; sum += wsum[k]; }
	*total = sum;
	return wbase + incl - v;
}

__global__ void __launch_bounds__( TR_THREADS, 8 ) k_tris_ordered( rc_scene_t sc, rc_work_t w, int32_t ntris )
{
	/* a translucent brush entity's scanline spans and a batch of triangle set-ups are never live together */
	union tr_shared_t
	{
		struct { rc_span_t spans[TR_THREADS / 32][32 * RC_TB_LINESPANS]; int32_t n[TR_THREADS / 32][32]; } tb;
		struct
		{
			rc_trisetup_t ts[TR_BATCH];
			int32_t y0[TR_BATCH], rows[TR_BATCH];
			int4 item[TR_BATCH];               /* tex_ofs, drawer, color, alpha of the triangle's item */
			rc_trirow_t row[TR_THREADS];
			int32_t pre[TR_THREADS + 1];
			uint32_t hit[TR_THREADS * 6];      /* polygon * 8 + fan triangle, in drawing order */
			uint8_t hslot[TR_THREADS * 6];     /* its item's slot in the chunk */
		} tr;
	};
	__shared__ tr_shared_t S;
	__shared__ int32_t s_wsum[TR_THREADS / 32];
	/* a chunk of the view's ordered items: item, first polygon (prefix sum of the triangle counts), polygon base */
	__shared__ int32_t s_iidx[TR_THREADS], s_ipre[TR_THREADS + 1], s_ibase[TR_THREADS];
	__shared__ uint32_t s_brush[TR_THREADS / 32];
	const int view = blockIdx.x, strip = blockIdx.y, t = threadIdx.x, lane = t & 31, warp = t >> 5;
	const int32_t ys0 = (int32_t)((long long)w.height * strip / gridDim.y ), ys1 = (int32_t)((long long)w.height * ( strip + 1 ) / gridDim.y );
	uint16_t *color = w.color + (size_t)view * w.width * w.height;
	int16_t *zb = w.zbuf + (size_t)view * w.width * w.height;
	const int32_t *order = w.item_order + (size_t)view * w.nitems;
	const rc_poly_t *polys = w.polys + (size_t)view * ntris;
	rc_blendluts_t L;
	L.colormap = sc.colormap; L.alphamap = w.alphamap; L.addmap = w.addmap; L.modmap = w.modmap;
	unsigned long long npix = 0;

	for( int32_t oc = 0; oc < w.nitems; oc += TR_THREADS )
	{
		/* thread = item of the chunk (the list ends with -1s) */
		const int32_t oi = oc + t;
		const int32_t idx = oi < w.nitems ? order[oi] : -1;
		int32_t cnt = 0, pbase = 0;
		bool brush = false;
		if( idx >= w.nfast_items )
		{
			const rc_item_t *it = &w.items[idx];
			if( it->tex_ofs >= 0 )
			{
				if( it->kind == RC_ITEM_BRUSH ) brush = true;
				else if( w.itemvis[(size_t)view * w.nitems + idx] ) { cnt = it->count; pbase = it->order0 - w.fast_tris; }
			}
		}
		int32_t npolys;
		const int32_t ipre = rc_block_excl_scan( cnt, s_wsum, t, &npolys ); /* its first barrier: the previous chunk is done */
		s_iidx[t] = idx; s_ipre[t] = ipre; s_ibase[t] = pbase;
		if( t == 0 ) s_ipre[TR_THREADS] = npolys;
		const uint32_t bm = __ballot_sync( 0xFFFFFFFFu, brush );
		if( lane == 0 ) s_brush[warp] = bm;
		if( __syncthreads_and( idx < 0 )) /* nothing but the list's tail */
			break;

		/* runs of items between translucent brush entities: the run's polygons, flat */
		for( int32_t ra = 0; ra < TR_THREADS; )
		{
			int32_t rb_ = TR_THREADS;
			for( int32_t wq = ra >> 5; wq < TR_THREADS / 32; wq++ )
			{
				uint32_t m = s_brush[wq];
				if( wq == ( ra >> 5 )) m &= 0xFFFFFFFFu << ( ra & 31 );
				if( m ) { rb_ = wq * 32 + __ffs( m ) - 1; break; }
			}
			const int32_t q0 = s_ipre[ra], q1 = s_ipre[rb_];
			for( int32_t qb = q0; qb < q1; qb += TR_THREADS )
			{
				/* thread = polygon: which of its fan triangles reach into the strip (the scanlines a triangle covers are
				 * [lowest v, highest v) of its corners: rc_tri_rows) */
				uint32_t hm = 0;
				int32_t slot = 0, pabs = 0;
				if( qb + t < q1 )
				{
					const int32_t q = qb + t;
					int32_t lo = ra, hi = rb_ - 1;
					while( lo < hi )
					{
						const int32_t mid = ( lo + hi + 1 ) >> 1;
						if( s_ipre[mid] <= q ) lo = mid; else hi = mid - 1;
					}
					slot = lo;
					pabs = s_ibase[slot] + ( q - s_ipre[slot] );
					const rc_poly_t *pl = &polys[pabs];
					const int32_t k = pl->k;
					if( k >= 3 )
					{
						const int32_t v0 = pl->v[0][1];
						int32_t va = pl->v[1][1];
						for( int32_t i = 1; i < k - 1; i++ )
						{
							const int32_t vb = pl->v[i + 1][1];
							const int32_t vlo = min( v0, min( va, vb )), vhi = max( v0, max( va, vb ));
							if( vlo < ys1 && vhi > ys0 ) hm |= 1u << i;
							va = vb;
						}
					}
				}
				int32_t nhit;
				int32_t at = rc_block_excl_scan( __popc( hm ), s_wsum, t, &nhit );
				if( nhit == 0 )
					continue;
				for( uint32_t m = hm; m; m &= m - 1, at++ )
				{
					S.tr.hit[at] = ((uint32_t)pabs << 3 ) | (uint32_t)( __ffs( m ) - 1 );
					S.tr.hslot[at] = (uint8_t)slot;
				}
				for( int32_t hb = 0; hb < nhit; hb += TR_BATCH )
				{
					__syncthreads(); /* the hit list is written; the previous batch's set-ups are free */
					const int32_t nb = min( TR_BATCH, nhit - hb );
					/* thread = triangle of the batch: the edge walk set-up (spread over the warps) */
					const int32_t bj = lane * ( TR_THREADS / 32 ) + warp;
					if( lane < TR_BATCH / ( TR_THREADS / 32 ) && bj < nb )
					{
						const uint32_t h = S.tr.hit[hb + bj];
						const rc_item_t *it = &w.items[s_iidx[S.tr.hslot[hb + bj]]];
						const rc_poly_t *pl = &polys[h >> 3];
						const int32_t i = (int32_t)( h & 7 );
						rc_fv_t a, b, c;
						const int32_t *pa = pl->v[0], *pb = pl->v[i], *pc = pl->v[i + 1];
						a.u = pa[0]; a.v = pa[1]; a.s = pa[2]; a.t = pa[3]; a.l = pa[4]; a.zi = pa[5];
						b.u = pb[0]; b.v = pb[1]; b.s = pb[2]; b.t = pb[3]; b.l = pb[4]; b.zi = pb[5];
						c.u = pc[0]; c.v = pc[1]; c.s = pc[2]; c.t = pc[3]; c.l = pc[4]; c.zi = pc[5];
						int32_t y0 = 0, rows = 0;
						if( rc_tri_setup( &a, &b, &c, it->skinwidth, it->skinheight, &S.tr.ts[bj] ))
							rows = rc_tri_rows( &S.tr.ts[bj], &y0 );
						S.tr.y0[bj] = y0; S.tr.rows[bj] = rows;
						int32_t icolor = it->color, ialpha = it->alpha;
						if( it->glow >= 0 )
						{
							const rc_glow_t *g = &w.glow[(size_t)view * w.nglow + it->glow];
							icolor = g->color; ialpha = g->alpha;
						}
						S.tr.item[bj] = make_int4( it->tex_ofs, it->drawer, icolor, ialpha );
					}
					__syncthreads();
					for( int32_t j = 0; j < nb; j++ )
					{
						const int32_t y0 = S.tr.y0[j];
						const int32_t rlo = ys0 > y0 ? ys0 - y0 : 0, rhi = min( S.tr.rows[j], ys1 - y0 );
						if( rlo >= rhi )
							continue;
						const rc_trisetup_t *ts = &S.tr.ts[j];
						const int4 iv = S.tr.item[j];
						const uint16_t *skin = sc.texels + iv.x;
						/* the triangle's pixels inside the strip, flat over the block: thread = scanline sets the rows up
						 * (closed form), a block prefix sum numbers the pixels, then thread = pixel.  A triangle writes a
						 * pixel once. */
						for( int32_t rb = rlo; rb < rhi; rb += TR_THREADS )
						{
							rc_trirow_t row;
							int32_t pc = 0;
							if( rb + t < rhi )
							{
								rc_tri_row( ts, rb + t, &row );
								if( row.count > 0 && rc_tri_row_inbounds( ts, &row ))
									pc = row.count;
							}
							int32_t total;
							const int32_t pre = rc_block_excl_scan( pc, s_wsum, t, &total ); /* its first barrier: the previous triangle's pixels are in the frame */
							if( total == 0 )
								continue;
							S.tr.row[t] = row;
							S.tr.pre[t] = pre;
							if( t == 0 ) S.tr.pre[TR_THREADS] = total;
							__syncthreads();
							for( int32_t p = t; p < total; p += TR_THREADS )
							{
								/* the row whose pixel range holds p */
								int32_t lo = 0, hi = TR_THREADS - 1;
								while( lo < hi )
								{
									const int32_t mid = ( lo + hi + 1 ) >> 1;
									if( S.tr.pre[mid] <= p ) lo = mid; else hi = mid - 1;
								}
								const rc_trirow_t &rw = S.tr.row[lo];
								const int32_t px = p - S.tr.pre[lo];
								const size_t at2 = (size_t)( y0 + rb + lo ) * w.width + rw.x + px;
								const uint32_t src = __ldg( skin + rc_tri_texel( ts, &rw, px ));
								const int32_t light = (int32_t)((uint32_t)rw.l + (uint32_t)px * (uint32_t)ts->lstepx );
								const int32_t zi = (int32_t)((uint32_t)rw.zi + (uint32_t)px * (uint32_t)ts->zistepx );
								rc_tri_shade( &L, iv.y, (uint32_t)iv.z, (uint32_t)iv.w, src, light, zi, &color[at2], &zb[at2] );
							}
							if( t == 0 ) npix += (unsigned long long)total;
							__syncthreads(); /* the rows are free; the pixels are in the frame */
						}
					}
				}
				__syncthreads(); /* the hit list is free */
			}
			if( rb_ < TR_THREADS )
			{
				__syncthreads(); /* the run's pixels are in the frame */
				rc_tbrush_draw( sc, w, &L, view, w.items[s_iidx[rb_]].first, ys0, ys1, color, zb, S.tb.spans[warp], S.tb.n[warp], lane, warp );
				__syncthreads();
			}
			ra = rb_ + 1;
		}
	}
	if( t == 0 && npix ) atomicAdd( &w.counters[RC_CTR_TRI_PIXELS], npix );
}

/* ---------------------------------------------------------------------------------------------- */
/*
 * The opaque fast path.  R_PolysetFillSpans8 (r_polyse.c:1134) is "if( z >= *pz ) { *pz = z; *pdest = shade; }" applied
 * triangle after triangle: the z a pixel ends with is the largest of the world's and every covering triangle's, and its
 * colour is that of the LAST triangle in submission order that reaches that largest z (an equal z overwrites).  Both are
 * the maximum of the 64-bit key (z << 32 | submission number): a run of opaque triangles needs no ordering at all -
 *   k_tri_keys_init  key = the world's z, number 0
 *   k_tris_fast      thread per (view, triangle): final verts, clip, edge walk; atomicMax of its key on every covered
 *                    pixel (read first: most fragments of a hidden model lose without an atomic); triangles with many
 *                    pixels are queued for k_tris_fast_big, a warp per triangle, lanes across the pixels of a scanline
 *   k_tri_resolve    thread per pixel whose key carries a triangle number: that triangle is set up again and this one
 *                    pixel shaded from it (texel, light level through the colour map), z from the key
 * The host sends the leading run of opaque items down this path (C4: every studio model); whatever follows it goes
 * through k_tris_ordered.
 */
#define TF_BIG_PIXELS 192

__global__ void __launch_bounds__( 256 ) k_tri_keys_init( rc_work_t w )
{
	const size_t n = (size_t)w.nviews * w.width * w.height;
	for( size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x )
		w.trikeys[i] = (unsigned long long)(uint32_t)( (int32_t)w.zbuf[i] + 32768 ) << 32;
}

/* one scanline of a set-up triangle into the key buffer; lanes: p0 = first pixel this caller takes, pstep = stride */
__device__ __forceinline__ uint32_t rc_fast_row( const rc_trisetup_t *ts, int32_t r, unsigned long long *krow, uint32_t id, int32_t p0, int32_t pstep )
{
	rc_trirow_t row;
	rc_tri_row( ts, r, &row );
	if( row.count <= 0 || !rc_tri_row_inbounds( ts, &row ))
		return 0;
	unsigned long long *kp = krow + row.x;
	for( int32_t p = p0; p < row.count; p += pstep )
	{
		const int32_t zi = (int32_t)((uint32_t)row.zi + (uint32_t)p * (uint32_t)ts->zistepx );
		const unsigned long long key = ((unsigned long long)(uint32_t)(( zi >> 16 ) + 32768 ) << 32 ) | id;
		if( key > kp[p] )
			atomicMax( &kp[p], key );
	}
	return (uint32_t)row.count;
}

#define TF_THREADS 128
#define TF_ROUNDS  24    /* triangles per thread: a block takes up to 3072 triangles of one (view, item) pair */
#define TF_QUEUE   256   /* fan triangles waiting for their edge walk */

struct tf_entry_t
{
	int32_t v[3][6];     /* u, v, s, t, l, zi of the three corners, in submission order */
	uint32_t id;         /* triangle * 8 + fan + 1 */
}; /* 76 B */

/* the edge walk of one queued triangle into the key buffer (or onto the big list) */
__device__ __forceinline__ void rc_fast_walk( const rc_work_t &w, const rc_item_t *it, int view, unsigned long long *keys, const tf_entry_t *e,
	unsigned *npix, unsigned *ntri )
{
	rc_fv_t a, b, c;
	a.u = e->v[0][0]; a.v = e->v[0][1]; a.s = e->v[0][2]; a.t = e->v[0][3]; a.l = e->v[0][4]; a.zi = e->v[0][5];
	b.u = e->v[1][0]; b.v = e->v[1][1]; b.s = e->v[1][2]; b.t = e->v[1][3]; b.l = e->v[1][4]; b.zi = e->v[1][5];
	c.u = e->v[2][0]; c.v = e->v[2][1]; c.s = e->v[2][2]; c.t = e->v[2][3]; c.l = e->v[2][4]; c.zi = e->v[2][5];
	rc_trisetup_t ts;
	if( !rc_tri_setup( &a, &b, &c, it->skinwidth, it->skinheight, &ts ))
		return;
	( *ntri )++;
	int32_t y0;
	const int32_t rows = rc_tri_rows( &ts, &y0 );
	/* a rough pixel count: the bounding box of the three corners */
	const int32_t umin = min( a.u, min( b.u, c.u )), umax = max( a.u, max( b.u, c.u ));
	if( rows * ( umax - umin + 1 ) > TF_BIG_PIXELS )
	{
		const int32_t slot = atomicAdd( w.big_count, 1 );
		if( slot < w.big_cap )
		{
			w.big_list[2 * slot] = view; w.big_list[2 * slot + 1] = (int32_t)e->id;
			return;
		}
	}
	for( int32_t r = 0; r < rows; r++ )
		*npix += rc_fast_row( &ts, r, keys + (size_t)( y0 + r ) * w.width, e->id, 0, 1 );
}

/*
 * Block per (view, item) pair and run of TF_THREADS * TF_ROUNDS triangles.  Most triangles of a model die early (back
 * faces, off screen): the threads project and clip 128 triangles at a time and queue the fan triangles that face the
 * viewer; whenever 256 are waiting the block walks their edges, one thread per triangle with every lane busy - the edge
 * walk is the long part, and run straight after the projection it kept 6 lanes of 32 busy (profiles/r02l).
 */
__global__ void __launch_bounds__( TF_THREADS, 8 ) k_tris_fast( rc_work_t w )
{
	__shared__ tf_entry_t s_q[TF_QUEUE];
	__shared__ int32_t s_qn;
	if( (int32_t)blockIdx.x >= *w.pair_count )
		return;
	const int32_t pair = w.pairs[blockIdx.x];
	const int view = pair >> 16, t = threadIdx.x;
	const rc_item_t *it = &w.items[pair & 0xFFFF];
	const int32_t first = blockIdx.y * ( TF_THREADS * TF_ROUNDS );
	if( first >= it->count )
		return;
	const int32_t last = min( it->count, first + TF_THREADS * TF_ROUNDS );
	const rc_view_t *vw = &w.views[view];
	unsigned long long *keys = w.trikeys + (size_t)view * w.width * w.height;
	unsigned npix = 0, ntri = 0;
	if( t == 0 ) s_qn = 0;
	__syncthreads();
	for( int32_t base = first; base < last; base += TF_THREADS )
	{
		const int32_t tri = base + t;
		if( tri < last )
		{
			rc_corner_t c[3];
			rc_fv_t poly[8];
			int k = 0;
			if( rc_item_corners( w, vw, it, tri, c ))
				k = rc_view_triangle( vw, &c[0], &c[1], &c[2], poly );
			for( int i = 1; i < k - 1; i++ )
			{
				/* the checks R_DrawTriangle starts with (r_polyse.c:636): degenerate, facing away */
				const int32_t abu = poly[0].u - poly[i].u, abv = poly[0].v - poly[i].v;
				const int32_t acu = poly[0].u - poly[i + 1].u, acv = poly[0].v - poly[i + 1].v;
				if( !( abu | abv ) || !( acu | acv ) || !(( acu * abv ) - ( abu * acv ) < 0 ))
					continue;
				tf_entry_t e;
				const rc_fv_t *src[3] = { &poly[0], &poly[i], &poly[i + 1] };
#pragma unroll
				for( int q = 0; q < 3; q++ )
				{
					e.v[q][0] = src[q]->u; e.v[q][1] = src[q]->v; e.v[q][2] = src[q]->s; e.v[q][3] = src[q]->t; e.v[q][4] = src[q]->l; e.v[q][5] = src[q]->zi;
				}
				e.id = (uint32_t)( it->order0 + tri ) * 8u + (uint32_t)i + 1u;
				const int32_t slot = atomicAdd( &s_qn, 1 );
				if( slot < TF_QUEUE ) s_q[slot] = e;
				else rc_fast_walk( w, it, view, keys, &e, &npix, &ntri ); /* a round of heavily clipped triangles: no room */
			}
		}
		__syncthreads();
		const int32_t n = min( s_qn, TF_QUEUE );
		const bool end = base + TF_THREADS >= last;
		__syncthreads(); /* every thread has read the count before the next round adds to it */
		if( n >= TF_THREADS || end )
		{
			/* whole passes of 128 triangles (at the end: what is left) */
			const int32_t take = end ? n : ( n / TF_THREADS ) * TF_THREADS;
			for( int32_t e = t; e < take; e += TF_THREADS )
				rc_fast_walk( w, it, view, keys, &s_q[e], &npix, &ntri );
			__syncthreads();
			/* the rest moves to the front */
			tf_entry_t keep;
			const bool mine = take + t < n;
			if( mine ) keep = s_q[take + t];
			__syncthreads();
			if( mine ) s_q[t] = keep;
			if( t == 0 ) s_qn = n - take;
			__syncthreads();
		}
	}
	/* counters: one atomic per warp */
	npix = __reduce_add_sync( 0xFFFFFFFFu, npix ); ntri = __reduce_add_sync( 0xFFFFFFFFu, ntri );
	if(( threadIdx.x & 31 ) == 0 )
	{
		if( npix ) atomicAdd( &w.counters[RC_CTR_TRI_PIXELS], (unsigned long long)npix );
		if( ntri ) atomicAdd( &w.counters[RC_CTR_TRIS], (unsigned long long)ntri );
	}
}

/* the set-up of fan triangle `fan` of triangle g for a view, again (k_tris_fast_big, k_tri_resolve); 0 = not drawn */
__device__ __forceinline__ int rc_fast_setup( const rc_work_t &w, int view, int32_t g, int32_t fan, rc_trisetup_t *ts, const rc_item_t **pit )
{
	const rc_item_t *it = rc_item_of( w, g );
	*pit = it;
	rc_corner_t c[3];
	rc_fv_t poly[8];
	if( !rc_item_corners( w, &w.views[view], it, g - it->order0, c ))
		return 0;
	const int k = rc_view_triangle( &w.views[view], &c[0], &c[1], &c[2], poly );
	if( fan < 1 || fan >= k - 1 )
		return 0;
	return rc_tri_setup( &poly[0], &poly[fan], &poly[fan + 1], it->skinwidth, it->skinheight, ts );
}

__global__ void __launch_bounds__( 128 ) k_tris_fast_big( rc_work_t w )
{
	const int lane = threadIdx.x & 31;
	int32_t n = *w.big_count;
	if( n > w.big_cap ) n = w.big_cap;
	if( blockIdx.x == 0 && threadIdx.x == 0 && n ) atomicAdd( &w.counters[RC_CTR_BIG_TRIS], (unsigned long long)n );
	const int32_t nwarps = gridDim.x * ( blockDim.x >> 5 );
	unsigned long long npix = 0;
	for( int32_t e = blockIdx.x * ( blockDim.x >> 5 ) + ( threadIdx.x >> 5 ); e < n; e += nwarps )
	{
		const int view = w.big_list[2 * e];
		const uint32_t id = (uint32_t)w.big_list[2 * e + 1];
		rc_trisetup_t ts;
		const rc_item_t *it;
		if( !rc_fast_setup( w, view, (int32_t)(( id - 1u ) >> 3 ), (int32_t)(( id - 1u ) & 7u ), &ts, &it ))
			continue;
		int32_t y0;
		const int32_t rows = rc_tri_rows( &ts, &y0 );
		unsigned long long *keys = w.trikeys + (size_t)view * w.width * w.height;
		/* lane per scanline: the edge stepping of a row costs more than its pixels */
		for( int32_t r = lane; r < rows; r += 32 )
			npix += rc_fast_row( &ts, r, keys + (size_t)( y0 + r ) * w.width, id, 0, 1 );
	}
	npix = __reduce_add_sync( 0xFFFFFFFFu, (unsigned)npix );
	if( lane == 0 && npix ) atomicAdd( &w.counters[RC_CTR_TRI_PIXELS], npix );
}

__global__ void __launch_bounds__( 128 ) k_tri_resolve( rc_scene_t sc, rc_work_t w )
{
	const size_t vpx = (size_t)w.width * w.height, n = (size_t)w.nviews * vpx;
	unsigned nw = 0;
	for( size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x )
	{
		const unsigned long long key = w.trikeys[i];
		const uint32_t id = (uint32_t)key;
		if( !id )
			continue;
		const int view = (int)( i / vpx );
		const int32_t pix = (int32_t)( i - (size_t)view * vpx ), y = pix / w.width, x = pix - y * w.width;
		rc_trisetup_t ts;
		const rc_item_t *it;
		if( !rc_fast_setup( w, view, (int32_t)(( id - 1u ) >> 3 ), (int32_t)(( id - 1u ) & 7u ), &ts, &it ))
			continue;
		int32_t y0;
		rc_tri_rows( &ts, &y0 );
		rc_trirow_t row;
		rc_tri_row( &ts, y - y0, &row );
		const int32_t p = x - row.x;
		const uint32_t src = __ldg( sc.texels + it->tex_ofs + rc_tri_texel( &ts, &row, p ));
		const uint32_t light = (uint32_t)row.l + (uint32_t)p * (uint32_t)ts.lstepx;
		w.color[i] = (uint16_t)( __ldg( sc.colormap + (( src >> 3 ) | (( light & 0x1F00u ) << 5 ))) | ( src & 7u ));
		w.zbuf[i] = (int16_t)( (int32_t)( key >> 32 ) - 32768 );
		nw++;
	}
	nw = __reduce_add_sync( 0xFFFFFFFFu, nw );
	if(( threadIdx.x & 31 ) == 0 && nw ) atomicAdd( &w.counters[RC_CTR_TRI_WRITTEN], (unsigned long long)nw );
}

/* ---------------------------------------------------------------------------------------------- */
/* the HUD pass: one launch per command (they overlap and their order is the result), thread per destination pixel */
__global__ void __launch_bounds__( 256 ) k_2d( rc_scene_t sc, rc_work_t w, const rc_cmd2d_t *cmds, int which, int view )
{
	const rc_cmd2d_t c = cmds[which];
	const rc_blendluts_t L = { sc.colormap, w.alphamap, w.addmap, w.modmap };
	const int32_t n = c.height * c.w;
	uint16_t *frame = w.color + (size_t)view * w.width * w.height;
	for( int32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x )
	{
		const int32_t v = i / c.w, u = i - v * c.w;
		rc_2d_pixel( &L, &c, sc.texels, v, (uint32_t)u, frame + (size_t)( c.y + v ) * w.width + c.x + u );
	}
}

#endif /* RC_K_TRIS_CUH */
