/*
 * rc_k_draw.cuh - S3b, span drawing: D_DrawSpans16 / D_DrawZSpans / Turbulent8 / D_BackgroundSurf
 * (ref/soft/r_scan.c:89-221,528-668,1309; r_edge.c:823) from the span lists k_spans left in HBM.
 *
 * One warp draws one scanline: lanes set up the 8-pixel (16 for turbulent) perspective chunks of the line's
 * spans into a shared-memory table (IEEE divides once per chunk, not per pixel), then walk the line in ALIGNED
 * groups of 8 pixels: texel gathers from the surface cache, one 16-byte colour store and one 16-byte z store
 * per lane - fully coalesced 512-byte rows per warp instruction.
 */
#ifndef RC_K_DRAW_CUH
#define RC_K_DRAW_CUH

#define DR_WARPS  4   /* scanlines per block */
#define SC_SPCAP  128 /* spans per draw batch */

struct rc_spaninfo_t
{
	int32_t u, end;      /* pixel range [u, end) */
	int32_t chunk0;      /* first entry of the span in the chunk table */
	int32_t kind;        /* RC_KIND_* */
	int32_t base, width; /* texel offset + row pitch of the cache block / texture; sky: base = colour */
	int32_t izi0, izistep;
}; /* 32 B */

static inline size_t rc_draw_smem( int32_t chunkcap )
{
	return (size_t)DR_WARPS * ( sizeof( rc_spaninfo_t ) * SC_SPCAP + sizeof( int4 ) * (size_t)chunkcap );
}

/* one pixel of a span: colour + z */
__device__ __forceinline__ void rc_shade( const rc_spaninfo_t &si, const int4 *tab, const uint16_t *cache, const uint16_t *texels,
	const int32_t *turb, int32_t x, uint32_t &col, uint32_t &z )
{
	const int32_t d = x - si.u;
	z = (uint32_t)((int32_t)((uint32_t)si.izi0 + (uint32_t)d * (uint32_t)si.izistep ) >> 16 ) & 0xFFFFu;
	if( si.kind == RC_KIND_SOLID )
	{
		const int4 c = tab[si.chunk0 + ( d >> 3 )];
		const uint32_t i = (uint32_t)( d & 7 );
		const int32_t s = (int32_t)((uint32_t)c.x + i * (uint32_t)c.z );
		const int32_t t = (int32_t)((uint32_t)c.y + i * (uint32_t)c.w );
		col = cache[si.base + ( s >> 16 ) + ( t >> 16 ) * si.width];
	}
	else if( si.kind == RC_KIND_TURB )
	{
		const int4 c = tab[si.chunk0 + ( d >> 4 )];
		const uint32_t i = (uint32_t)( d & 15 );
		const int32_t s = (int32_t)((uint32_t)c.x + i * (uint32_t)c.z );
		const int32_t t = (int32_t)((uint32_t)c.y + i * (uint32_t)c.w );
		const int32_t sturb = (( s + turb[( t >> 16 ) & ( RC_CYCLE - 1 )] ) >> 16 ) & 63;
		const int32_t tturb = (( t + turb[( s >> 16 ) & ( RC_CYCLE - 1 )] ) >> 16 ) & 63;
		col = texels[si.base + ( tturb << 6 ) + sturb];
	}
	else if( si.kind == RC_KIND_SKY )
		col = (uint32_t)si.base & 0xFFFFu;
	else
	{
		col = 0; z = 0; /* background under clear */
	}
}

/* draw phase: one warp draws the spans of one scanline */
__device__ __noinline__ void rc_draw_line( const rc_scene_t &sc, const rc_work_t &w, const rc_view_t *vw, const rc_surf_t *vsurfs,
	const rc_span_t *spans, int32_t nsp, int32_t y, rc_spaninfo_t *info, int4 *tab, uint16_t *crow, int16_t *zrow, int lane )
{
	const int32_t *turb = sc.sintable + vw->turb_phase;
	const bool vec_ok = ( w.width & 7 ) == 0;

	for( int32_t b0 = 0; b0 < nsp; b0 += SC_SPCAP )
	{
		const int32_t nb = min( SC_SPCAP, nsp - b0 );

		/* D0: per-span records + first chunk of every span */
		int32_t total = 0;
		for( int32_t base = 0; base < nb; base += 32 )
		{
			const int32_t i = base + lane;
			int32_t c = 0;
			rc_spaninfo_t si;
			if( i < nb )
			{
				const rc_span_t sp = spans[b0 + i];
				si.u = sp.u; si.end = sp.u + sp.count;
				si.kind = RC_KIND_BACKGROUND; si.base = 0; si.width = 0; si.izi0 = 0; si.izistep = 0;
				if( sp.surf )
				{
					const rc_surf_t *sf = &vsurfs[sp.surf - 1];
					si.kind = rc_span_kind( sf );
					si.base = si.kind == RC_KIND_SKY ? vw->sky_color : sf->cache_ofs;
					si.width = sf->cachewidth;
					rc_span_z0( sf, sp.u, y, &si.izi0, &si.izistep );
					if( si.kind != RC_KIND_SKY )
						c = rc_span_nchunks( si.kind, sp.count );
				}
			}
			const int32_t x = rc_warp_incl_scan( c, lane );
			if( i < nb )
			{
				si.chunk0 = total + x - c;
				info[i] = si;
			}
			total += __shfl_sync( 0xFFFFFFFFu, x, 31 );
		}
		__syncwarp();

		/* D1: chunk table (s, t, sstep, tstep at the first pixel of every chunk) */
		for( int32_t c = lane; c < total; c += 32 )
		{
			int32_t lo = 0, hi = nb - 1;
			while( lo < hi )
			{
				const int32_t mid = ( lo + hi + 1 ) >> 1;
				if( info[mid].chunk0 <= c ) lo = mid; else hi = mid - 1;
			}
			const rc_span_t sp = spans[b0 + lo];
			int4 r;
			rc_chunk_st( &vsurfs[sp.surf - 1], info[lo].kind, sp.u, y, sp.count, c - info[lo].chunk0, &r.x, &r.y, &r.z, &r.w );
			tab[c] = r;
		}
		__syncwarp();

		/* D2: aligned groups of 8 pixels */
		const int32_t xa = info[0].u, xb = info[nb - 1].end;
		for( int32_t g = ( xa >> 3 ) + lane; ( g << 3 ) < xb; g += 32 )
		{
			const int32_t X = g << 3;
			const int32_t xs = max( X, xa );
			int32_t lo = 0, hi = nb - 1;
			while( lo < hi )
			{
				const int32_t mid = ( lo + hi + 1 ) >> 1;
				if( info[mid].u <= xs ) lo = mid; else hi = mid - 1;
			}
			rc_spaninfo_t si = info[lo];
			uint32_t cw[4] = { 0, 0, 0, 0 }, zw[4] = { 0, 0, 0, 0 };
			uint32_t mask = 0;
#pragma unroll
			for( int k = 0; k < 8; k++ )
			{
				const int32_t x = X + k;
				if( x < xa || x >= xb )
					continue;
				while( x >= si.end )
					si = info[++lo];
				if( si.kind == RC_KIND_BACKGROUND && !w.clear )
					continue;
				uint32_t col, z;
				rc_shade( si, tab, sc.cache, sc.texels, turb, x, col, z );
				cw[k >> 1] |= col << (( k & 1 ) * 16 );
				zw[k >> 1] |= z << (( k & 1 ) * 16 );
				mask |= 1u << k;
			}
			if( mask == 0xFFu && vec_ok )
			{
				*reinterpret_cast<uint4 *>( crow + X ) = make_uint4( cw[0], cw[1], cw[2], cw[3] );
				*reinterpret_cast<uint4 *>( zrow + X ) = make_uint4( zw[0], zw[1], zw[2], zw[3] );
			}
			else
			{
#pragma unroll
				for( int k = 0; k < 8; k++ )
					if( mask & ( 1u << k ))
					{
						crow[X + k] = (uint16_t)( cw[k >> 1] >> (( k & 1 ) * 16 ));
						zrow[X + k] = (int16_t)( zw[k >> 1] >> (( k & 1 ) * 16 ));
					}
			}
		}
		__syncwarp();
	}
}



__global__ void __launch_bounds__( DR_WARPS * 32 ) k_draw( rc_scene_t sc, rc_work_t w, int32_t chunkcap )
{
	extern __shared__ __align__( 16 ) unsigned char smem_raw[];
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	rc_spaninfo_t *info = reinterpret_cast<rc_spaninfo_t *>( smem_raw ) + warp * SC_SPCAP;
	int4 *tab = reinterpret_cast<int4 *>( smem_raw + sizeof( rc_spaninfo_t ) * SC_SPCAP * DR_WARPS ) + (size_t)warp * chunkcap;

	const int view = blockIdx.y;
	const int y = blockIdx.x * DR_WARPS + warp;
	if( y >= w.height )
		return;
	const rc_lineinfo_t li = w.lineinfo[(size_t)view * w.height + y];
	if( li.nsp <= 0 )
		return;
	const rc_span_t *spans = reinterpret_cast<const rc_span_t *>( w.events + (size_t)view * w.view_events + li.ofs );
	rc_draw_line( sc, w, &w.views[view], w.surfs + (size_t)view * w.max_cands, spans, li.nsp, y, info, tab,
		w.color + ((size_t)view * w.height + y ) * w.width, w.zbuf + ((size_t)view * w.height + y ) * w.width, lane );
}

#endif /* RC_K_DRAW_CUH */
