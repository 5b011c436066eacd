/*
 * rc_k_draw.cuh - S3b, span drawing: D_DrawSpans16 / D_DrawZSpans / Turbulent8 / D_BackgroundSurf
 * (ref/soft/r_scan.c:89-221,528-668,1309; r_edge.c:823) from the span lists k_lines left in HBM.
 *
 * One warp draws one scanline at a time (lines strided over the grid), in batches of up to 32 spans:
 *   D0  lane = span: 1/z and s/z, t/z at the span's first pixel, its first texel (one IEEE divide), the integer
 *       1/z stepping of D_DrawZSpans, the number of 8-pixel (16 for turbulent) subdivisions; warp prefix sum ->
 *       every span's slice of the chunk table.  Span records go to shared memory.
 *   D1  lane = subdivision ("chunk"): the reference accumulates s/z, t/z, 1/z in float from chunk to chunk, so chunk
 *       j repeats those j additions, then ONE divide gives its first texel; the texel the reference calls snext is
 *       the first texel of chunk j + 1 = the neighbour lane's value (shuffle), so only the last chunk of a span (and
 *       lane 31) pays a second divide.  {s, t, sstep, tstep} go to the chunk table in shared memory.
 *   D2  span by span (warp-uniform span data), lane = pixel: 16.16 texel address from the chunk entry, one texel
 *       gather from the surface cache, 1/z by integer stepping; colour and z leave as contiguous 64-byte rows per
 *       warp instruction.
 */
#ifndef RC_K_DRAW_CUH
#define RC_K_DRAW_CUH

#define DR_WARPS  4   /* warps per block */
#define SC_SPCAP  32  /* spans per draw batch = lanes */
#define DR_GRID   ( 148 * 16 )

struct rc_spaninfo_t
{
	int32_t u, end;      /* pixel range [u, end) */
	int32_t chunk0;      /* first entry of the span in the chunk table */
	int32_t kind;        /* RC_KIND_* */
	int32_t base, width; /* texel offset + row pitch of the cache block / texture; sky: base = colour */
	int32_t izi0, izistep;
}; /* 32 B */

struct rc_spanext_t
{
	float   sdivz, tdivz, zi; /* at the span's first pixel */
	int32_t s0, t0;           /* its first texel (16.16) */
	int32_t surf;             /* candidate index */
	int32_t count, pad;
}; /* 32 B */

static inline int32_t rc_draw_chunkcap( int32_t width ) { return ( width + 7 ) / 8 + SC_SPCAP + 1; }

static inline size_t rc_draw_smem( int32_t chunkcap )
{
	return (size_t)DR_WARPS * (( sizeof( rc_spaninfo_t ) + sizeof( rc_spanext_t )) * SC_SPCAP + sizeof( int4 ) * (size_t)chunkcap );
}

/* clamped 16.16 texel coordinate at (sdivz, zi): r_scan.c:559-571 */
__device__ __forceinline__ void rc_texel_at( const rc_surf_t *sf, float sdivz, float tdivz, float zi, int32_t lo, int32_t *s, int32_t *t )
{
	const float z = (float)0x10000 / zi;
	*s = rc_clamp_st( rc_f2i( sdivz * z ) + sf->sadjust, sf->bbextents, lo );
	*t = rc_clamp_st( rc_f2i( tdivz * z ) + sf->tadjust, sf->bbextentt, lo );
}

__global__ void __launch_bounds__( DR_WARPS * 32 ) k_draw( rc_scene_t sc, rc_work_t w, int32_t chunkcap, int32_t line0, int32_t line1 )
{
	extern __shared__ __align__( 16 ) unsigned char smem_raw[];
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	rc_spaninfo_t *info = reinterpret_cast<rc_spaninfo_t *>( smem_raw ) + warp * SC_SPCAP;
	rc_spanext_t *ext = reinterpret_cast<rc_spanext_t *>( smem_raw + sizeof( rc_spaninfo_t ) * SC_SPCAP * DR_WARPS ) + warp * SC_SPCAP;
	int4 *tab = reinterpret_cast<int4 *>( smem_raw + ( sizeof( rc_spaninfo_t ) + sizeof( rc_spanext_t )) * SC_SPCAP * DR_WARPS ) + (size_t)warp * chunkcap;

	const uint16_t *__restrict__ cache = sc.cache;
	const uint16_t *__restrict__ texels = sc.texels;
	const int32_t nwarps = gridDim.x * DR_WARPS; /* lines [line0, line1) of the batch: the host pipelines ranges of views against their copies out */
	const bool clear = w.clear != 0;

	for( int32_t line = line0 + blockIdx.x * DR_WARPS + warp; line < line1; line += nwarps )
	{
		const rc_lineinfo_t li = w.lineinfo[line];
		if( li.nsp <= 0 )
			continue;
		const int32_t view = line / w.height, y = line - view * w.height;
		const rc_span_t *spans = reinterpret_cast<const rc_span_t *>( w.events + (size_t)view * w.view_events + li.ofs );
		const rc_surf_t *vsurfs = w.surfs + (size_t)view * w.max_cands;
		const rc_view_t *vw = &w.views[view];
		const int32_t *turb = sc.sintable + vw->turb_phase;
		uint16_t *crow = w.color + (size_t)line * w.width;
		int16_t *zrow = w.zbuf + (size_t)line * w.width;
		const float dv = (float)y;

		for( int32_t b0 = 0; b0 < li.nsp; b0 += SC_SPCAP )
		{
			const int32_t nb = min( SC_SPCAP, li.nsp - b0 );

			/* D0: per-span records */
			int32_t c = 0;
			rc_spaninfo_t si;
			si.u = 0; si.end = 0; si.kind = RC_KIND_BACKGROUND; si.base = 0; si.width = 0; si.izi0 = 0; si.izistep = 0;
			if( lane < nb )
			{
				const rc_span_t sp = spans[b0 + lane];
				si.u = sp.u; si.end = sp.u + sp.count;
				rc_spanext_t ex;
				ex.sdivz = ex.tdivz = ex.zi = 0; ex.s0 = ex.t0 = 0; ex.surf = sp.surf - 1; ex.count = sp.count; ex.pad = 0;
				if( sp.surf )
				{
					const rc_surf_t *sf = &vsurfs[sp.surf - 1];
					si.kind = rc_span_kind( sf );
					si.base = si.kind == RC_KIND_SKY ? vw->sky_color : sf->cache_ofs;
					si.width = sf->cachewidth;
					rc_span_z0( sf, sp.u, y, &si.izi0, &si.izistep );
					if( si.kind != RC_KIND_SKY )
					{
						c = rc_span_nchunks( si.kind, sp.count );
						const float du = (float)sp.u;
						ex.zi = sf->d_ziorigin + dv * sf->d_zistepv + du * sf->d_zistepu;
						ex.sdivz = sf->d_sdivzorigin + dv * sf->d_sdivzstepv + du * sf->d_sdivzstepu;
						ex.tdivz = sf->d_tdivzorigin + dv * sf->d_tdivzstepv + du * sf->d_tdivzstepu;
						rc_texel_at( sf, ex.sdivz, ex.tdivz, ex.zi, 0, &ex.s0, &ex.t0 );
					}
				}
				ext[lane] = ex;
			}
			const int32_t incl = rc_warp_incl_scan( c, lane );
			si.chunk0 = incl - c;
			info[lane] = si;
			const int32_t total = __shfl_sync( 0xFFFFFFFFu, incl, 31 );
			__syncwarp();

			/* D1: chunk table (s, t, sstep, tstep at the first pixel of every chunk) */
			for( int32_t cb = 0; cb < total; cb += 32 )
			{
				const int32_t ci = cb + lane;
				const bool live = ci < total;
				int32_t s_raw = 0, t_raw = 0, j = 0, nch = 1, px = 8, sidx = 0;
				float sdivz = 0, tdivz = 0, zi = 1.0f;
				const rc_surf_t *sf = vsurfs;
				bool turbk = false;
				if( live )
				{
					int32_t lo = 0, hi = nb - 1;
					while( lo < hi )
					{
						const int32_t mid = ( lo + hi + 1 ) >> 1;
						if( info[mid].chunk0 <= ci ) lo = mid; else hi = mid - 1;
					}
					sidx = lo;
					const rc_spanext_t ex = ext[lo];
					sf = &vsurfs[ex.surf];
					turbk = info[lo].kind == RC_KIND_TURB;
					px = turbk ? 16 : 8;
					j = ci - info[lo].chunk0;
					nch = rc_span_nchunks( info[lo].kind, ex.count );
					const float fpx = (float)px;
					const float sdivzstep = sf->d_sdivzstepu * fpx, tdivzstep = sf->d_tdivzstepu * fpx, zistep = sf->d_zistepu * fpx;
					sdivz = ex.sdivz; tdivz = ex.tdivz; zi = ex.zi;
					for( int32_t i = 0; i < j; i++ )
					{
						sdivz += sdivzstep; tdivz += tdivzstep; zi += zistep;
					}
					s_raw = ex.s0; t_raw = ex.t0;
					if( j )
						rc_texel_at( sf, sdivz, tdivz, zi, px, &s_raw, &t_raw );
				}
				/* snext / tnext: the neighbour lane holds chunk j + 1 of the same span unless this is the span's last chunk */
				int32_t snext = __shfl_down_sync( 0xFFFFFFFFu, s_raw, 1 );
				int32_t tnext = __shfl_down_sync( 0xFFFFFFFFu, t_raw, 1 );
				if( live )
				{
					const rc_spanext_t ex = ext[sidx];
					const int32_t spancount = ex.count - j * px;
					const bool last = j == nch - 1;
					const bool own = last ? spancount > 1 : ( lane == 31 || ci + 1 >= total );
					if( own )
					{
						float a, b, d;
						if( last )
						{
							const float spancountminus1 = (float)( spancount - 1 );
							a = sdivz + sf->d_sdivzstepu * spancountminus1;
							b = tdivz + sf->d_tdivzstepu * spancountminus1;
							d = zi + sf->d_zistepu * spancountminus1;
						}
						else
						{
							const float fpx = (float)px;
							a = sdivz + sf->d_sdivzstepu * fpx;
							b = tdivz + sf->d_tdivzstepu * fpx;
							d = zi + sf->d_zistepu * fpx;
						}
						rc_texel_at( sf, a, b, d, px, &snext, &tnext );
					}
					/* D_DrawSpans16 never reloads s from snext: every later chunk start is congruent to s0 modulo 8
					 * (r_scan.c:576-607); Turbulent8 does reload */
					int32_t s = s_raw, t = t_raw;
					if( j && !turbk )
					{
						s -= ( s - ex.s0 ) & 7;
						t -= ( t - ex.t0 ) & 7;
					}
					int32_t sstep = 0, tstep = 0;
					if( !last )
					{
						const int sh = turbk ? 4 : 3;
						sstep = ( snext - s ) >> sh;
						tstep = ( tnext - t ) >> sh;
					}
					else if( spancount > 1 )
					{
						sstep = ( snext - s ) / ( spancount - 1 );
						tstep = ( tnext - t ) / ( spancount - 1 );
					}
					if( turbk )
					{
						s &= ( RC_CYCLE << 16 ) - 1;
						t &= ( RC_CYCLE << 16 ) - 1;
					}
					tab[ci] = make_int4( s, t, sstep, tstep );
				}
			}
			__syncwarp();

			/* D2: span by span, lane = pixel */
			for( int32_t k = 0; k < nb; k++ )
			{
				const rc_spaninfo_t sp = info[k];
				if( sp.kind == RC_KIND_BACKGROUND && !clear )
					continue;
				const int4 *stab = tab + sp.chunk0;
				if( sp.kind == RC_KIND_SOLID )
				{
					const uint16_t *blk = cache + sp.base;
					for( int32_t x = sp.u + lane; x < sp.end; x += 32 )
					{
						const uint32_t d = (uint32_t)( x - sp.u );
						const int4 cc = stab[d >> 3];
						const uint32_t i = d & 7u;
						const int32_t s = (int32_t)((uint32_t)cc.x + i * (uint32_t)cc.z );
						const int32_t t = (int32_t)((uint32_t)cc.y + i * (uint32_t)cc.w );
						const uint16_t col = __ldg( blk + ( s >> 16 ) + ( t >> 16 ) * sp.width );
						crow[x] = col;
						zrow[x] = (int16_t)((int32_t)((uint32_t)sp.izi0 + d * (uint32_t)sp.izistep ) >> 16 );
					}
				}
				else if( sp.kind == RC_KIND_TURB )
				{
					const uint16_t *tex = texels + sp.base;
					for( int32_t x = sp.u + lane; x < sp.end; x += 32 )
					{
						const uint32_t d = (uint32_t)( x - sp.u );
						const int4 cc = stab[d >> 4];
						const uint32_t i = d & 15u;
						const int32_t s = (int32_t)((uint32_t)cc.x + i * (uint32_t)cc.z );
						const int32_t t = (int32_t)((uint32_t)cc.y + i * (uint32_t)cc.w );
						const int32_t sturb = (( s + turb[( t >> 16 ) & ( RC_CYCLE - 1 )] ) >> 16 ) & 63;
						const int32_t tturb = (( t + turb[( s >> 16 ) & ( RC_CYCLE - 1 )] ) >> 16 ) & 63;
						crow[x] = __ldg( tex + ( tturb << 6 ) + sturb );
						zrow[x] = (int16_t)((int32_t)((uint32_t)sp.izi0 + d * (uint32_t)sp.izistep ) >> 16 );
					}
				}
				else if( sp.kind == RC_KIND_SKY )
				{
					for( int32_t x = sp.u + lane; x < sp.end; x += 32 )
					{
						const uint32_t d = (uint32_t)( x - sp.u );
						crow[x] = (uint16_t)sp.base;
						zrow[x] = (int16_t)((int32_t)((uint32_t)sp.izi0 + d * (uint32_t)sp.izistep ) >> 16 );
					}
				}
				else
				{
					for( int32_t x = sp.u + lane; x < sp.end; x += 32 )
					{
						crow[x] = 0; zrow[x] = 0; /* background under clear */
					}
				}
			}
			__syncwarp();
		}
	}
}

#endif /* RC_K_DRAW_CUH */
