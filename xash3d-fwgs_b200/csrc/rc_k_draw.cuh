/*
 * rc_k_draw.cuh - S3b, span drawing: D_DrawSpans16 / D_DrawZSpans / Turbulent8 / D_BackgroundSurf
 * (ref/soft/r_scan.c:89-221,528-668,1309; r_edge.c:823) from the span lists k_lines left in HBM.
 *
 * One warp draws a group of lpw consecutive scanlines of a view at a time (groups strided over the grid; lpw = 2 for
 * narrow frames, where one line has too few spans and subdivisions to fill 32 lanes - 320 pixels: 45 subdivisions = 2
 * rounds at 70 %, two lines: 3 rounds at 94 %).  The group's spans are one list, in batches of up to 32:
 *   D0  lane = span: 1/z and s/z, t/z at the span's first pixel, its first texel (one IEEE divide), the integer
 *       1/z stepping of D_DrawZSpans, the per-subdivision float steps, the number of 8-pixel (16 for turbulent)
 *       subdivisions; warp prefix sum -> every span's range of subdivision numbers.  Records go to shared memory.
 *   D1  lane = subdivision ("chunk"): the reference accumulates s/z, t/z, 1/z in float from chunk to chunk, so chunk
 *       j repeats those j additions, then ONE divide gives its first texel; the texel the reference calls snext is
 *       the first texel of chunk j + 1 = the neighbour lane's value (shuffle), so only the last chunk of a span (and
 *       lane 31) pays a second divide.  The same lane then walks its 8 (16) pixels with the reference's integer
 *       stepping - texel gather from the surface cache, 1/z step - into a scanline buffer in shared memory.
 *   D2  (before D1: the pixel ranges are disjoint) sky / background spans (no subdivisions) fill their part of the
 *       buffer; after the last batch the group leaves as aligned 16-byte colour and z stores.
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

/* lines a warp draws together */
static inline int32_t rc_draw_lpw( int32_t width, int32_t height )
{
	return ( width <= 1024 && !( height & 1 )) ? 2 : 1;
}

/* shared memory per warp: span records + lpw scanlines of colour and of z (pitch = frame width: the lines of a group
 * are contiguous, in the buffer as in the frame) */
static inline size_t rc_draw_smem( int32_t width, int32_t lpw )
{
	const size_t w8 = ((size_t)width * lpw + 7 ) & ~(size_t)7;
	return (size_t)DR_WARPS * (( sizeof( rc_spaninfo_t ) + sizeof( rc_spanext_t )) * SC_SPCAP + 4 * w8 );
}

/* buffer range [xa, xb) -> frame rows: aligned 16-byte stores for the body, 2-byte stores for a ragged head / tail */
__device__ __forceinline__ void rc_rows_out( uint16_t *crow, uint16_t *zrow, const uint16_t *cbuf, const uint16_t *zbuf,
	int32_t xa, int32_t xb, bool vec_ok, int lane )
{
	if( xa >= xb )
		return;
	const int32_t a8 = vec_ok ? ( xa + 7 ) & ~7 : xb, b8 = vec_ok ? xb & ~7 : xb;
	if( a8 >= b8 )
	{
		for( int32_t x = xa + lane; x < xb; x += 32 ) { crow[x] = cbuf[x]; zrow[x] = zbuf[x]; }
		return;
	}
	for( int32_t x = xa + lane; x < a8; x += 32 ) { crow[x] = cbuf[x]; zrow[x] = zbuf[x]; }
	for( int32_t x = a8 + lane * 8; x < b8; x += 256 )
	{
		*reinterpret_cast<uint4 *>( crow + x ) = *reinterpret_cast<const uint4 *>( cbuf + x );
		*reinterpret_cast<uint4 *>( zrow + x ) = *reinterpret_cast<const uint4 *>( zbuf + x );
	}
	for( int32_t x = b8 + lane; x < xb; x += 32 ) { crow[x] = cbuf[x]; zrow[x] = zbuf[x]; }
}

/* clamped 16.16 texel coordinate at (sdivz, zi): r_scan.c:559-571 */
__device__ __forceinline__ void rc_texel_at( const rc_spanext_t &ex, float sdivz, float tdivz, float zi, int32_t lo, int32_t *s, int32_t *t )
{
	const float z = (float)0x10000 / zi;
	*s = rc_clamp_st( rc_f2i( sdivz * z ) + ex.sadjust, ex.bbextents, lo );
	*t = rc_clamp_st( rc_f2i( tdivz * z ) + ex.tadjust, ex.bbextentt, lo );
}

/* n / d (C truncation) for 1 <= d <= 15 with m = ceil(2^31 / d): exact while |n| < 2^27 (e * |n| < 2^31 for the
 * magic number's excess e < d), which 16.16 texel differences inside a surface always are; the divide instruction
 * sequence (some 30 instructions, twice per last subdivision) stays as the fallback */
__device__ __forceinline__ int32_t rc_div_small( int32_t n, int32_t d, uint32_t m )
{
	const uint32_t a = n < 0 ? 0u - (uint32_t)n : (uint32_t)n;
	if( a >= ( 1u << 27 ))
		return n / d;
	const uint32_t q = (uint32_t)(((unsigned long long)a * m ) >> 31 );
	return n < 0 ? -(int32_t)q : (int32_t)q;
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

__global__ void __launch_bounds__( DR_WARPS * 32, 8 ) k_draw( rc_scene_t sc, rc_work_t w, int32_t line0, int32_t line1, int32_t lpw )
{
	extern __shared__ __align__( 16 ) unsigned char smem_raw[];
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const int32_t w8 = ( w.width * lpw + 7 ) & ~7;
	rc_spaninfo_t *info = reinterpret_cast<rc_spaninfo_t *>( smem_raw ) + warp * SC_SPCAP;
	rc_spanext_t *ext = reinterpret_cast<rc_spanext_t *>( smem_raw + sizeof( rc_spaninfo_t ) * SC_SPCAP * DR_WARPS ) + warp * SC_SPCAP;
	uint16_t *cbuf = reinterpret_cast<uint16_t *>( smem_raw + ( sizeof( rc_spaninfo_t ) + sizeof( rc_spanext_t )) * SC_SPCAP * DR_WARPS ) + (size_t)warp * 2 * w8;
	uint16_t *zbuf = cbuf + w8;

	const uint16_t *__restrict__ cache = sc.cache;
	const uint16_t *__restrict__ texels = sc.texels;
	const int32_t nwarps = gridDim.x * DR_WARPS; /* lines [line0, line1) of the batch: whole views (the host pipelines ranges of views against their copies out) */
	const bool clear = w.clear != 0;
	const bool vec_ok = ( w.width & 7 ) == 0;
	const int32_t ngroups = ( line1 - line0 ) / lpw; /* lpw divides the frame height */
	__shared__ uint32_t s_magic[16];
	if( threadIdx.x > 0 && threadIdx.x < 16 ) s_magic[threadIdx.x] = ( 0x80000000u + threadIdx.x - 1u ) / threadIdx.x;
	__syncthreads();

	for( int32_t g = blockIdx.x * DR_WARPS + warp; g < ngroups; g += nwarps )
	{
		const int32_t line = line0 + g * lpw; /* the group's first line */
		const int32_t view = line / w.height, y = line - view * w.height;
		/* the group's span list: line A's spans, then line B's */
		int32_t nA = 0, nB = 0;
		const rc_span_t *spA = nullptr, *spB = nullptr;
		{
			const rc_lineinfo_t li = w.lineinfo[line];
			if(( li.nsp & RC_LINE_DONE ) && li.ofs >= 0 )
			{
				nA = li.nsp & ~RC_LINE_DONE;
				spA = reinterpret_cast<const rc_span_t *>( w.events + (size_t)view * w.view_events + li.ofs );
			}
			if( lpw > 1 )
			{
				const rc_lineinfo_t lb = w.lineinfo[line + 1];
				if(( lb.nsp & RC_LINE_DONE ) && lb.ofs >= 0 )
				{
					nB = lb.nsp & ~RC_LINE_DONE;
					spB = reinterpret_cast<const rc_span_t *>( w.events + (size_t)view * w.view_events + lb.ofs );
				}
			}
		}
		const int32_t nsp = nA + nB;
		if( nsp <= 0 )
			continue;
		const rc_surf_t *vsurfs = w.surfs + (size_t)view * w.max_cands;
		const rc_view_t *vw = &w.views[view];
		const int32_t *turb = sc.sintable + vw->turb_phase;
		uint16_t *crow = w.color + (size_t)line * w.width;
		uint16_t *zrow = reinterpret_cast<uint16_t *>( w.zbuf ) + (size_t)line * w.width;
		/* buffer ranges the two lines' spans cover */
		int32_t xaA = 0x7FFFFFFF, xbA = 0, xaB = 0x7FFFFFFF, xbB = 0;

		for( int32_t b0 = 0; b0 < nsp; b0 += SC_SPCAP )
		{
			const int32_t nb = min( SC_SPCAP, nsp - b0 );

			/* D0: per-span records; a span of line B sits one frame row further in the buffer */
			int32_t c = 0, spu = 0, sy = y;
			rc_spaninfo_t si;
			rc_surf_t sf;
			si.u = 0; si.end = 0; si.kind = RC_KIND_BACKGROUND; si.base = 0; si.width = 0; si.izi0 = 0; si.izistep = 0;
			const bool isB = b0 + lane >= nA;
			if( lane < nb )
			{
				const rc_span_t sp = isB ? spB[b0 + lane - nA] : spA[b0 + lane];
				const int32_t rowofs = isB ? w.width : 0;
				sy = isB ? y + 1 : y; spu = sp.u;
				si.u = sp.u + rowofs; si.end = si.u + sp.count;
				if( sp.surf )
				{
					sf = rc_load_surf( &vsurfs[sp.surf - 1] );
					si.kind = rc_span_kind( &sf );
					si.base = si.kind == RC_KIND_SKY ? vw->sky_color : sf.cache_ofs;
					si.width = sf.cachewidth;
					rc_span_z0( &sf, sp.u, sy, &si.izi0, &si.izistep );
					if( si.kind != RC_KIND_SKY )
						c = rc_span_nchunks( si.kind, sp.count );
				}
			}
			/* records of the spans that have subdivisions, packed: record r = the r-th such span */
			const uint32_t cmask = __ballot_sync( 0xFFFFFFFFu, c > 0 );
			const int32_t incl = rc_warp_incl_scan( c, lane );
			const int32_t chunk0 = incl - c;
			si.chunk0 = chunk0;
			if( c > 0 )
			{
				const int32_t r = __popc( cmask & (( 1u << lane ) - 1u ));
				const float dv = (float)sy, du = (float)spu, fpx = si.kind == RC_KIND_TURB ? 16.0f : 8.0f;
				rc_spanext_t ex;
				ex.zi = sf.d_ziorigin + dv * sf.d_zistepv + du * sf.d_zistepu;
				ex.sdivz = sf.d_sdivzorigin + dv * sf.d_sdivzstepv + du * sf.d_sdivzstepu;
				ex.tdivz = sf.d_tdivzorigin + dv * sf.d_tdivzstepv + du * sf.d_tdivzstepu;
				ex.count = si.end - si.u;
				ex.sdivzstep = sf.d_sdivzstepu * fpx; ex.tdivzstep = sf.d_tdivzstepu * fpx; ex.zistep = sf.d_zistepu * fpx;
				ex.sdivzstepu = sf.d_sdivzstepu; ex.tdivzstepu = sf.d_tdivzstepu; ex.zistepu = sf.d_zistepu;
				ex.sadjust = sf.sadjust; ex.tadjust = sf.tadjust; ex.bbextents = sf.bbextents; ex.bbextentt = sf.bbextentt;
				rc_texel_at( ex, ex.sdivz, ex.tdivz, ex.zi, 0, &ex.s0, &ex.t0 );
				info[r] = si;
				ext[r] = ex;
			}
			const int32_t total = __shfl_sync( 0xFFFFFFFFu, incl, 31 );
			{
				const bool inA = lane < nb && !isB, inB = lane < nb && isB;
				xaA = min( xaA, (int32_t)__reduce_min_sync( 0xFFFFFFFFu, inA ? (uint32_t)si.u : 0x7FFFFFFFu ));
				xbA = max( xbA, (int32_t)__reduce_max_sync( 0xFFFFFFFFu, inA ? (uint32_t)si.end : 0u ));
				xaB = min( xaB, (int32_t)__reduce_min_sync( 0xFFFFFFFFu, inB ? (uint32_t)si.u : 0x7FFFFFFFu ));
				xbB = max( xbB, (int32_t)__reduce_max_sync( 0xFFFFFFFFu, inB ? (uint32_t)si.end : 0u ));
			}
			__syncwarp();

			/* D2: spans without subdivisions - sky (flat colour, z from the sky plane), background; their records are
			 * still in the registers of the lanes that set them up */
			for( uint32_t fm = __ballot_sync( 0xFFFFFFFFu, lane < nb && c == 0 ); fm; fm &= fm - 1 )
			{
				const int src = __ffs( (int)fm ) - 1;
				const int32_t fu = __shfl_sync( 0xFFFFFFFFu, si.u, src ), fend = __shfl_sync( 0xFFFFFFFFu, si.end, src );
				const int32_t fkind = __shfl_sync( 0xFFFFFFFFu, si.kind, src );
				if( fkind == RC_KIND_SKY )
				{
					const int32_t fbase = __shfl_sync( 0xFFFFFFFFu, si.base, src );
					const uint32_t fizi0 = (uint32_t)__shfl_sync( 0xFFFFFFFFu, si.izi0, src ), fstep = (uint32_t)__shfl_sync( 0xFFFFFFFFu, si.izistep, src );
					for( int32_t x = fu + lane; x < fend; x += 32 )
					{
						cbuf[x] = (uint16_t)fbase;
						zbuf[x] = (uint16_t)( (int32_t)( fizi0 + (uint32_t)( x - fu ) * fstep ) >> 16 );
					}
				}
				else if( fkind == RC_KIND_BACKGROUND )
				{
					/* ref_soft leaves these pixels alone (D_SolidSurf returns on the background surface); a cleared
					 * batch starts from zeros, otherwise the buffer keeps what is there */
					for( int32_t x = fu + lane; x < fend; x += 32 )
					{
						cbuf[x] = clear ? (uint16_t)0 : crow[x];
						zbuf[x] = clear ? (uint16_t)0 : zrow[x];
					}
				}
			}

			/* D1: lane = subdivision: its first texel + steps, then its pixels */
			for( int32_t cb = 0; cb < total; cb += 32 )
			{
				const int32_t ci = cb + lane;
				const bool live = ci < total;
				int32_t s_raw = 0, t_raw = 0, j = 0, px = 8, sidx = 0;
				float sdivz = 0, tdivz = 0, zi = 1.0f;
				bool turbk = false;
				/* the record of subdivision ci: spans that started before this round + span heads up to this lane */
				const uint32_t heads = __reduce_or_sync( 0xFFFFFFFFu, ( c > 0 && chunk0 >= cb && chunk0 < cb + 32 ) ? 1u << ( chunk0 - cb ) : 0u );
				const int32_t nbefore = __popc( __ballot_sync( 0xFFFFFFFFu, c > 0 && chunk0 < cb ));
				if( live )
				{
					const int32_t lo = nbefore + __popc( heads & ( 0xFFFFFFFFu >> ( 31 - lane ))) - 1;
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
						const uint32_t m = s_magic[spancount - 1];
						sstep = rc_div_small( snext - s, spancount - 1, m );
						tstep = rc_div_small( tnext - t, spancount - 1, m );
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

			__syncwarp();
		}

		/* the group leaves: one contiguous range when line A ends where line B starts (spans cover whole rows) */
		if( xbA == xaB )
			rc_rows_out( crow, zrow, cbuf, zbuf, xaA, xbB, vec_ok, lane );
		else
		{
			rc_rows_out( crow, zrow, cbuf, zbuf, xaA, xbA, vec_ok, lane );
			rc_rows_out( crow, zrow, cbuf, zbuf, xaB, xbB, vec_ok, lane );
		}
		__syncwarp();
	}
}

#endif /* RC_K_DRAW_CUH */
