/*
 * rc_k_draw.cuh - S3b, span drawing: D_DrawSpans16 / D_DrawZSpans / Turbulent8 / D_BackgroundSurf
 * (ref/soft/r_scan.c:89-221,528-668,1309; r_edge.c:823) from the span lists k_lines left in HBM.
 *
 * One warp draws one scanline at a time (lines strided over the grid), in batches of up to 32 spans:
 *   D0  lane = span: 1/z and s/z, t/z at the span's first pixel, its first texel (one IEEE divide), the integer
 *       1/z stepping of D_DrawZSpans, the per-subdivision float steps, the number of 8-pixel (16 for turbulent)
 *       subdivisions; warp prefix sum -> every span's range of subdivision numbers.  Records go to shared memory.
 *   D1  lane = subdivision ("chunk"): the reference accumulates s/z, t/z, 1/z in float from chunk to chunk, so chunk
 *       j repeats those j additions, then ONE divide gives its first texel; the texel the reference calls snext is
 *       the first texel of chunk j + 1 = the neighbour lane's value (shuffle), so only the last chunk of a span (and
 *       lane 31) pays a second divide.  The same lane then walks its 8 (16) pixels with the reference's integer
 *       stepping - texel gather from the surface cache, 1/z step - into a scanline buffer in shared memory.
 *   D2  sky / background spans (no subdivisions) fill their part of the buffer, then the scanline leaves as aligned
 *       16-byte colour and z stores.
 */
#ifndef RC_K_DRAW_CUH
#define RC_K_DRAW_CUH

#define DR_WARPS  4   /* warps per block */
#define SC_SPCAP  32  /* spans per draw batch = lanes */
#define DR_GRID   ( 148 * 16 )

struct rc_spaninfo_t
{
	int32_t u, end;      /* pixel range [u, end) */
	int32_t chunk0;      /* number of the span's first subdivision inside the batch */
	int32_t kind;        /* RC_KIND_* */
	int32_t base, width; /* texel offset + row pitch of the cache block / texture; sky: base = colour */
	int32_t izi0, izistep;
}; /* 32 B */

struct rc_spanext_t
{
	float   sdivz, tdivz, zi;             /* at the span's first pixel */
	int32_t s0, t0;                       /* its first texel (16.16) */
	int32_t count;
	float   sdivzstep, tdivzstep, zistep; /* per subdivision (d_*stepu * 8 or 16) */
	float   sdivzstepu, tdivzstepu, zistepu;
	int32_t sadjust, tadjust, bbextents, bbextentt;
}; /* 64 B */

/* shared memory per warp: span records + a scanline of colour and of z */
static inline size_t rc_draw_smem( int32_t width )
{
	const size_t w8 = ((size_t)width + 7 ) & ~(size_t)7;
	return (size_t)DR_WARPS * (( sizeof( rc_spaninfo_t ) + sizeof( rc_spanext_t )) * SC_SPCAP + 4 * w8 );
}

/* clamped 16.16 texel coordinate at (sdivz, zi): r_scan.c:559-571 */
__device__ __forceinline__ void rc_texel_at( const rc_spanext_t &ex, float sdivz, float tdivz, float zi, int32_t lo, int32_t *s, int32_t *t )
{
	const float z = (float)0x10000 / zi;
	*s = rc_clamp_st( rc_f2i( sdivz * z ) + ex.sadjust, ex.bbextents, lo );
	*t = rc_clamp_st( rc_f2i( tdivz * z ) + ex.tadjust, ex.bbextentt, lo );
}

/* rc_surf_t (80 B) as five 16-byte loads */
__device__ __forceinline__ rc_surf_t rc_load_surf( const rc_surf_t *p )
{
	union { rc_surf_t s; uint4 q[5]; } u;
	const uint4 *src = reinterpret_cast<const uint4 *>( p );
#pragma unroll
	for( int i = 0; i < 5; i++ ) u.q[i] = __ldg( src + i );
	return u.s;
}

static_assert( sizeof( rc_surf_t ) == 80, "rc_load_surf copies five 16-byte words" );

__global__ void __launch_bounds__( DR_WARPS * 32 ) k_draw( rc_scene_t sc, rc_work_t w, int32_t line0, int32_t line1 )
{
	extern __shared__ __align__( 16 ) unsigned char smem_raw[];
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const int32_t w8 = ( w.width + 7 ) & ~7;
	rc_spaninfo_t *info = reinterpret_cast<rc_spaninfo_t *>( smem_raw ) + warp * SC_SPCAP;
	rc_spanext_t *ext = reinterpret_cast<rc_spanext_t *>( smem_raw + sizeof( rc_spaninfo_t ) * SC_SPCAP * DR_WARPS ) + warp * SC_SPCAP;
	uint16_t *cbuf = reinterpret_cast<uint16_t *>( smem_raw + ( sizeof( rc_spaninfo_t ) + sizeof( rc_spanext_t )) * SC_SPCAP * DR_WARPS ) + (size_t)warp * 2 * w8;
	uint16_t *zbuf = cbuf + w8;

	const uint16_t *__restrict__ cache = sc.cache;
	const uint16_t *__restrict__ texels = sc.texels;
	const int32_t nwarps = gridDim.x * DR_WARPS; /* lines [line0, line1) of the batch: the host pipelines ranges of views against their copies out */
	const bool clear = w.clear != 0;
	const bool vec_ok = ( w.width & 7 ) == 0;

	for( int32_t line = line0 + blockIdx.x * DR_WARPS + warp; line < line1; line += nwarps )
	{
		rc_lineinfo_t li = w.lineinfo[line];
		if( !( li.nsp & RC_LINE_DONE ))
			continue;
		li.nsp &= ~RC_LINE_DONE;
		if( li.nsp <= 0 )
			continue;
		const int32_t view = line / w.height, y = line - view * w.height;
		const rc_span_t *spans = reinterpret_cast<const rc_span_t *>( w.events + (size_t)view * w.view_events + li.ofs );
		const rc_surf_t *vsurfs = w.surfs + (size_t)view * w.max_cands;
		const rc_view_t *vw = &w.views[view];
		const int32_t *turb = sc.sintable + vw->turb_phase;
		uint16_t *crow = w.color + (size_t)line * w.width;
		uint16_t *zrow = reinterpret_cast<uint16_t *>( w.zbuf ) + (size_t)line * w.width;
		const float dv = (float)y;
		int32_t xa = 0x7FFFFFFF, xb = 0; /* pixel range of the line's spans */

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
				if( sp.surf )
				{
					const rc_surf_t sf = rc_load_surf( &vsurfs[sp.surf - 1] );
					si.kind = rc_span_kind( &sf );
					si.base = si.kind == RC_KIND_SKY ? vw->sky_color : sf.cache_ofs;
					si.width = sf.cachewidth;
					rc_span_z0( &sf, sp.u, y, &si.izi0, &si.izistep );
					if( si.kind != RC_KIND_SKY )
					{
						c = rc_span_nchunks( si.kind, sp.count );
						const float du = (float)sp.u, fpx = si.kind == RC_KIND_TURB ? 16.0f : 8.0f;
						rc_spanext_t ex;
						ex.zi = sf.d_ziorigin + dv * sf.d_zistepv + du * sf.d_zistepu;
						ex.sdivz = sf.d_sdivzorigin + dv * sf.d_sdivzstepv + du * sf.d_sdivzstepu;
						ex.tdivz = sf.d_tdivzorigin + dv * sf.d_tdivzstepv + du * sf.d_tdivzstepu;
						ex.count = sp.count;
						ex.sdivzstep = sf.d_sdivzstepu * fpx; ex.tdivzstep = sf.d_tdivzstepu * fpx; ex.zistep = sf.d_zistepu * fpx;
						ex.sdivzstepu = sf.d_sdivzstepu; ex.tdivzstepu = sf.d_tdivzstepu; ex.zistepu = sf.d_zistepu;
						ex.sadjust = sf.sadjust; ex.tadjust = sf.tadjust; ex.bbextents = sf.bbextents; ex.bbextentt = sf.bbextentt;
						rc_texel_at( ex, ex.sdivz, ex.tdivz, ex.zi, 0, &ex.s0, &ex.t0 );
						ext[lane] = ex;
					}
				}
			}
			const int32_t incl = rc_warp_incl_scan( c, lane );
			si.chunk0 = incl - c;
			info[lane] = si;
			const int32_t total = __shfl_sync( 0xFFFFFFFFu, incl, 31 );
			{
				const int32_t ua = __shfl_sync( 0xFFFFFFFFu, si.u, 0 ), ub = __shfl_sync( 0xFFFFFFFFu, si.end, nb - 1 );
				xa = min( xa, ua ); xb = max( xb, ub );
			}
			__syncwarp();

			/* D1: lane = subdivision: its first texel + steps, then its pixels */
			for( int32_t cb = 0; cb < total; cb += 32 )
			{
				const int32_t ci = cb + lane;
				const bool live = ci < total;
				int32_t s_raw = 0, t_raw = 0, j = 0, px = 8, sidx = 0;
				float sdivz = 0, tdivz = 0, zi = 1.0f;
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
					turbk = info[lo].kind == RC_KIND_TURB;
					px = turbk ? 16 : 8;
					j = ci - info[lo].chunk0;
					const rc_spanext_t &ex = ext[lo];
					const float sdivzstep = ex.sdivzstep, tdivzstep = ex.tdivzstep, zistep = ex.zistep;
					sdivz = ex.sdivz; tdivz = ex.tdivz; zi = ex.zi;
					for( int32_t i = 0; i < j; i++ )
					{
						sdivz += sdivzstep; tdivz += tdivzstep; zi += zistep;
					}
					s_raw = ex.s0; t_raw = ex.t0;
					if( j )
						rc_texel_at( ex, sdivz, tdivz, zi, px, &s_raw, &t_raw );
				}
				/* snext / tnext: the neighbour lane holds chunk j + 1 of the same span unless this is the span's last chunk */
				int32_t snext = __shfl_down_sync( 0xFFFFFFFFu, s_raw, 1 );
				int32_t tnext = __shfl_down_sync( 0xFFFFFFFFu, t_raw, 1 );
				if( live )
				{
					const rc_spanext_t &ex = ext[sidx];
					const rc_spaninfo_t sp = info[sidx];
					const int32_t spancount = ex.count - j * px;
					const bool last = spancount <= px;
					const bool own = last ? spancount > 1 : ( lane == 31 || ci + 1 >= total );
					if( own )
					{
						float a, b, d;
						if( last )
						{
							const float spancountminus1 = (float)( spancount - 1 );
							a = sdivz + ex.sdivzstepu * spancountminus1;
							b = tdivz + ex.tdivzstepu * spancountminus1;
							d = zi + ex.zistepu * spancountminus1;
						}
						else
						{
							a = sdivz + ex.sdivzstep;
							b = tdivz + ex.tdivzstep;
							d = zi + ex.zistep;
						}
						rc_texel_at( ex, a, b, d, px, &snext, &tnext );
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

					/* the chunk's pixels: integer stepping only (r_scan.c:600-640, 160-200, 1330-1350) */
					const int32_t x0 = sp.u + j * px;
					const int32_t npx = last ? spancount : px;
					uint32_t izi = (uint32_t)sp.izi0 + (uint32_t)( j * px ) * (uint32_t)sp.izistep;
					if( !turbk )
					{
						const uint16_t *blk = cache + sp.base;
#pragma unroll
						for( int i = 0; i < 8; i++ )
							if( i < npx )
							{
								cbuf[x0 + i] = __ldg( blk + ( s >> 16 ) + ( t >> 16 ) * sp.width );
								zbuf[x0 + i] = (uint16_t)( (int32_t)izi >> 16 );
								s = (int32_t)((uint32_t)s + (uint32_t)sstep ); t = (int32_t)((uint32_t)t + (uint32_t)tstep );
								izi += (uint32_t)sp.izistep;
							}
					}
					else
					{
						const uint16_t *tex = texels + sp.base;
						s &= ( RC_CYCLE << 16 ) - 1;
						t &= ( RC_CYCLE << 16 ) - 1;
						for( int i = 0; i < npx; i++ )
						{
							const int32_t sturb = (( s + turb[( t >> 16 ) & ( RC_CYCLE - 1 )] ) >> 16 ) & 63;
							const int32_t tturb = (( t + turb[( s >> 16 ) & ( RC_CYCLE - 1 )] ) >> 16 ) & 63;
							cbuf[x0 + i] = __ldg( tex + ( tturb << 6 ) + sturb );
							zbuf[x0 + i] = (uint16_t)( (int32_t)izi >> 16 );
							s = (int32_t)((uint32_t)s + (uint32_t)sstep ); t = (int32_t)((uint32_t)t + (uint32_t)tstep );
							izi += (uint32_t)sp.izistep;
						}
					}
				}
			}

			/* D2: spans without subdivisions - sky (flat colour, z from the sky plane), background */
			for( int32_t k = 0; k < nb; k++ )
			{
				const rc_spaninfo_t sp = info[k];
				if( sp.kind == RC_KIND_SKY )
				{
					for( int32_t x = sp.u + lane; x < sp.end; x += 32 )
					{
						cbuf[x] = (uint16_t)sp.base;
						zbuf[x] = (uint16_t)( (int32_t)((uint32_t)sp.izi0 + (uint32_t)( x - sp.u ) * (uint32_t)sp.izistep ) >> 16 );
					}
				}
				else if( sp.kind == RC_KIND_BACKGROUND )
				{
					/* ref_soft leaves these pixels alone (D_SolidSurf returns on the background surface); a cleared
					 * batch starts from zeros, otherwise the buffer keeps what is there */
					for( int32_t x = sp.u + lane; x < sp.end; x += 32 )
					{
						cbuf[x] = clear ? (uint16_t)0 : crow[x];
						zbuf[x] = clear ? (uint16_t)0 : zrow[x];
					}
				}
			}
			__syncwarp();
		}

		/* the scanline leaves: aligned 16-byte stores for the body, 2-byte stores for a ragged head / tail */
		if( xa < xb )
		{
			const int32_t a8 = vec_ok ? ( xa + 7 ) & ~7 : xb, b8 = vec_ok ? xb & ~7 : xb;
			if( a8 >= b8 )
			{
				for( int32_t x = xa + lane; x < xb; x += 32 ) { crow[x] = cbuf[x]; zrow[x] = zbuf[x]; }
			}
			else
			{
				for( int32_t x = xa + lane; x < a8; x += 32 ) { crow[x] = cbuf[x]; zrow[x] = zbuf[x]; }
				for( int32_t x = a8 + lane * 8; x < b8; x += 256 )
				{
					*reinterpret_cast<uint4 *>( crow + x ) = *reinterpret_cast<const uint4 *>( cbuf + x );
					*reinterpret_cast<uint4 *>( zrow + x ) = *reinterpret_cast<const uint4 *>( zbuf + x );
				}
				for( int32_t x = b8 + lane; x < xb; x += 32 ) { crow[x] = cbuf[x]; zrow[x] = zbuf[x]; }
			}
		}
		__syncwarp();
	}
}

#endif /* RC_K_DRAW_CUH */
