/*
 * rc_k_draw.cuh - S3b, span drawing: D_DrawSpans16 / D_DrawZSpans / Turbulent8 / D_BackgroundSurf
 * (ref/soft/r_scan.c:89-221,528-668,1309; r_edge.c:823) from the span lists k_lines left in HBM.
 *
 * One warp draws a group of DR_LINES consecutive scanlines of a view at a time (groups strided over the grid).  The
 * group's spans form one list, taken in batches of 32; the unit of pixel work is the OCTET, 8 consecutive pixels of a
 * span counted from its first pixel (a D_DrawSpans16 subdivision; half a Turbulent8 subdivision):
 *   D0  lane = span: 1/z and s/z, t/z at the span's first pixel, its first texel (one IEEE divide), the integer 1/z
 *       stepping of D_DrawZSpans, the number of octets; warp prefix sum -> every span's range of octet numbers.
 *   D1  lane = octet, 32 per round: the reference accumulates s/z, t/z, 1/z in float from subdivision to subdivision, so
 *       the lane repeats those additions - from the span start, or from the state the previous round's last lane left
 *       (at most 31 of them) - then ONE divide gives its first texel; the texel the reference calls snext is the first
 *       texel of the next subdivision = the neighbour lane's value (shuffle), so only the last subdivision of a span
 *       (and lane 31) pays a second divide.  The lane leaves a 32-byte record in shared memory: 16.16 texel coordinates
 *       of the octet's first pixel, their per-pixel steps, 1/z and its step, where the pixels go, where the texels are.
 *   D2  lane = 4 pixels of an octet (two lanes per octet, 16 octets per pass): integer stepping only - texel gather from
 *       the surface cache (the water texture, the sky colour), 1/z - and the pixels leave straight for the frame: the
 *       two lanes of an octet cover 16 contiguous bytes of the colour row and 16 of the z row, neighbouring octets of a
 *       span continue them, so the warp's stores are runs of whole 32-byte sectors; nothing is staged in shared memory
 *       and nothing is written twice.
 */
#ifndef RC_K_DRAW_CUH
#define RC_K_DRAW_CUH

#define DR_WARPS  4   /* warps per block */
#define SC_SPCAP  32  /* spans per draw batch = lanes */
#define DR_GRID   ( 148 * 16 )

struct rc_spaninfo_t
{
	int32_t x0;          /* first pixel, counted from the first pixel of the group's first line */
	int32_t count;
	int32_t oct0;        /* number of the span's first octet inside the batch */
	int32_t kind;        /* RC_KIND_* */
	int32_t base, width; /* texel offset + row pitch of the cache block / texture; sky: base = colour */
	int32_t izi0, izistep;
}; /* 32 B */

struct rc_spanext_t
{
	float   sdivz, tdivz, zi;             /* at the span's first pixel */
	int32_t s0, t0;                       /* its first texel (16.16) */
	float   sdivzstepu, tdivzstepu, zistepu;
	int32_t sadjust, tadjust, bbextents, bbextentt;
}; /* 48 B */

/* what D2 needs of an octet */
struct rc_octet_t
{
	int32_t s, t, sstep, tstep; /* 16.16 texel coordinates of the first pixel, per-pixel steps */
	int32_t izi, izistep;       /* 1/z (<< 31) of the first pixel, per-pixel step */
	int32_t base;               /* texel offset of the cache block / texture; sky: the colour */
	uint32_t where;             /* x0 (13 bits) | pixels (4) << 13 | kind (2) << 17 | row pitch of the texels (13) << 19 */
}; /* 32 B */

/* lines a warp draws together: enough of them that a batch of 32 spans is full (a 320-pixel line has ~10 spans) */
static inline int32_t rc_draw_lpw( int32_t width, int32_t height )
{
	int32_t g = width <= 512 ? 4 : ( width <= 1024 ? 2 : 1 );
	while( g > 1 && ( height % g )) g >>= 1;
	return g;
}

/* octets a batch of 32 spans can have: the pixels of the group's lines / 8 + one partial octet per span */
static inline int32_t rc_draw_chaincap( int32_t width, int32_t lpw )
{
	return ( width * lpw + 7 ) / 8 + SC_SPCAP;
}

static inline size_t rc_draw_smem( int32_t width, int32_t lpw )
{
	return (size_t)DR_WARPS * (( sizeof( rc_spaninfo_t ) + sizeof( rc_spanext_t )) * SC_SPCAP + sizeof( rc_octet_t ) * 32
		+ 3 * sizeof( float ) * (size_t)rc_draw_chaincap( width, lpw ));
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
static_assert( sizeof( rc_octet_t ) == 32, "D2 reads an octet record as two 16-byte words" );

__global__ void __launch_bounds__( DR_WARPS * 32, 8 ) k_draw( rc_scene_t sc, rc_work_t w, int32_t line0, int32_t line1, int32_t lpw )
{
	extern __shared__ __align__( 16 ) unsigned char smem_raw[];
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	rc_spaninfo_t *info = reinterpret_cast<rc_spaninfo_t *>( smem_raw ) + warp * SC_SPCAP;
	rc_spanext_t *ext = reinterpret_cast<rc_spanext_t *>( smem_raw + sizeof( rc_spaninfo_t ) * SC_SPCAP * DR_WARPS ) + warp * SC_SPCAP;
	rc_octet_t *octs = reinterpret_cast<rc_octet_t *>( smem_raw + ( sizeof( rc_spaninfo_t ) + sizeof( rc_spanext_t )) * SC_SPCAP * DR_WARPS ) + warp * 32;
	const int32_t chaincap = ( w.width * lpw + 7 ) / 8 + SC_SPCAP;
	/* s/z, t/z, 1/z at the first pixel of every octet of the batch (the reference's float accumulation, walked per span) */
	float *chain = reinterpret_cast<float *>( smem_raw + (( sizeof( rc_spaninfo_t ) + sizeof( rc_spanext_t )) * SC_SPCAP + sizeof( rc_octet_t ) * 32 ) * DR_WARPS ) + (size_t)warp * 3 * chaincap;

	const uint16_t *__restrict__ cache = sc.cache;
	const uint16_t *__restrict__ texels = sc.texels;
	const int32_t nwarps = gridDim.x * DR_WARPS; /* lines [line0, line1) of the batch: whole views */
	const bool clear = w.clear != 0;
	const int32_t ngroups = ( line1 - line0 ) / lpw; /* lpw divides the frame height */
	__shared__ uint32_t s_magic[16];
	if( threadIdx.x > 0 && threadIdx.x < 16 ) s_magic[threadIdx.x] = ( 0x80000000u + threadIdx.x - 1u ) / threadIdx.x;
	__syncthreads();

	for( int32_t g = blockIdx.x * DR_WARPS + warp; g < ngroups; g += nwarps )
	{
		const int32_t line = line0 + g * lpw; /* the group's first line */
		const int32_t view = line / w.height, y = line - view * w.height;
		/* the group's span list: the spans of its lines one after the other; lane l < lpw holds line l's slice */
		int32_t my_n = 0;
		const rc_span_t *my_sp = nullptr;
		if( lane < lpw )
		{
			const rc_lineinfo_t li = w.lineinfo[line + lane];
			if(( li.nsp & RC_LINE_DONE ) && li.ofs >= 0 )
			{
				my_n = li.nsp & ~RC_LINE_DONE;
				my_sp = reinterpret_cast<const rc_span_t *>( w.events + (size_t)view * w.view_events + li.ofs );
			}
		}
		const int32_t my_end = rc_warp_incl_scan( my_n, lane ); /* spans up to and including this lane's line */
		const int32_t nsp = __shfl_sync( 0xFFFFFFFFu, my_end, 31 );
		if( nsp <= 0 )
			continue;
		const rc_surf_t *vsurfs = w.surfs + (size_t)view * w.max_cands;
		const rc_view_t *vw = &w.views[view];
		const int32_t *turb = sc.sintable + vw->turb_phase;
		const bool texfilt = vw->texfilt != 0;
		uint16_t *crow = w.color + (size_t)line * w.width;
		uint16_t *zrow = reinterpret_cast<uint16_t *>( w.zbuf ) + (size_t)line * w.width;

		for( int32_t b0 = 0; b0 < nsp; b0 += SC_SPCAP )
		{
			const int32_t nb = min( SC_SPCAP, nsp - b0 );

			/* D0: per-span records.  Span b0 + lane belongs to the first line whose running total exceeds it */
			int32_t c = 0;
			{
				const int32_t si_ = b0 + lane;
				int32_t ln = 0;
				for( int l = 0; l < lpw; l++ )
					ln += __shfl_sync( 0xFFFFFFFFu, my_end, l ) <= si_ ? 1 : 0;
				const int32_t prev_end = __shfl_sync( 0xFFFFFFFFu, my_end, ln > 0 ? ln - 1 : 0 ); /* every lane takes part */
				const int32_t lstart = ln ? prev_end : 0;
				const unsigned long long lp = __shfl_sync( 0xFFFFFFFFu, (unsigned long long)(size_t)my_sp, ln < lpw ? ln : 0 );
				if( lane < nb )
				{
					const rc_span_t sp = reinterpret_cast<const rc_span_t *>( (size_t)lp )[si_ - lstart];
					rc_spaninfo_t si;
					si.x0 = sp.u + ln * w.width; si.count = sp.count;
					si.kind = RC_KIND_BACKGROUND; si.base = 0; si.width = 0; si.izi0 = 0; si.izistep = 0;
					c = ( sp.count + 7 ) >> 3;
					if( sp.surf )
					{
						const rc_surf_t sf = rc_load_surf( &vsurfs[sp.surf - 1] );
						const int32_t sy = y + ln;
						si.kind = rc_span_kind( &sf );
						si.base = si.kind == RC_KIND_SKY ? vw->sky_color : sf.cache_ofs;
						si.width = sf.cachewidth;
						rc_span_z0( &sf, sp.u, sy, &si.izi0, &si.izistep );
						if( si.kind != RC_KIND_SKY )
						{
							const float dv = (float)sy, du = (float)sp.u;
							rc_spanext_t ex;
							ex.zi = sf.d_ziorigin + dv * sf.d_zistepv + du * sf.d_zistepu;
							ex.sdivz = sf.d_sdivzorigin + dv * sf.d_sdivzstepv + du * sf.d_sdivzstepu;
							ex.tdivz = sf.d_tdivzorigin + dv * sf.d_tdivzstepv + du * sf.d_tdivzstepu;
							ex.sdivzstepu = sf.d_sdivzstepu; ex.tdivzstepu = sf.d_tdivzstepu; ex.zistepu = sf.d_zistepu;
							ex.sadjust = sf.sadjust; ex.tadjust = sf.tadjust; ex.bbextents = sf.bbextents; ex.bbextentt = sf.bbextentt;
							rc_texel_at( ex, ex.sdivz, ex.tdivz, ex.zi, 0, &ex.s0, &ex.t0 );
							ext[lane] = ex;
						}
					}
					else if( !clear )
						c = 0; /* ref_soft leaves background pixels alone (D_SolidSurf returns on the background surface) */
					info[lane] = si;
				}
			}
			const int32_t incl = rc_warp_incl_scan( c, lane );
			const int32_t oct0 = incl - c;
			if( lane < nb ) info[lane].oct0 = oct0;
			const int32_t total = __shfl_sync( 0xFFFFFFFFu, incl, 31 );
			/* the reference accumulates s/z, t/z, 1/z in float from subdivision to subdivision: the span's lane walks that
			 * chain once and leaves the state at the head of every octet (a 16-pixel subdivision fills two) */
			if( lane < nb && c > 0 )
			{
				const int32_t kd = info[lane].kind;
				if( kd == RC_KIND_SOLID || kd == RC_KIND_TURB )
				{
					const rc_spanext_t &ex = ext[lane];
					const bool turbk = kd == RC_KIND_TURB;
					const float fpx = turbk ? 16.0f : 8.0f;
					const float sdivzstep = ex.sdivzstepu * fpx, tdivzstep = ex.tdivzstepu * fpx, zistep = ex.zistepu * fpx;
					float a = ex.sdivz, b = ex.tdivz, d = ex.zi;
					float *dst = chain + 3 * oct0;
					if( !turbk )
						for( int32_t j = 0; j < c; j++, dst += 3 )
						{
							dst[0] = a; dst[1] = b; dst[2] = d;
							a += sdivzstep; b += tdivzstep; d += zistep;
						}
					else
						for( int32_t o2 = 0; o2 < c; o2++, dst += 3 )
						{
							dst[0] = a; dst[1] = b; dst[2] = d;
							if( o2 & 1 ) { a += sdivzstep; b += tdivzstep; d += zistep; }
						}
				}
			}
			__syncwarp();

			for( int32_t r0 = 0; r0 < total; r0 += 32 )
			{
				/* D1: lane = octet r0 + lane.  Its span: spans that started before this round + span heads up to this lane */
				const int32_t q = r0 + lane;
				const bool live = q < total;
				const uint32_t heads = __reduce_or_sync( 0xFFFFFFFFu, ( c > 0 && oct0 >= r0 && oct0 < r0 + 32 ) ? 1u << ( oct0 - r0 ) : 0u );
				const uint32_t before = __ballot_sync( 0xFFFFFFFFu, c > 0 && oct0 < r0 );
				/* spans without octets (background of an uncleared batch) have no record: index = the n-th span with octets */
				const uint32_t haveoct = __ballot_sync( 0xFFFFFFFFu, c > 0 );
				int32_t sidx = 0, o = 0, kind = RC_KIND_BACKGROUND, j = 0, px = 8;
				int32_t s_raw = 0, t_raw = 0;
				float sdivz = 0, tdivz = 0, zi = 1.0f;
				bool floaty = false;
				if( live )
				{
					const int32_t nth = __popc( before ) + __popc( heads & ( 0xFFFFFFFFu >> ( 31 - lane ))) - 1;
					sidx = nth;
					if( haveoct != ( nb >= 32 ? 0xFFFFFFFFu : ( 1u << nb ) - 1u ))
					{
						/* the nth set bit of haveoct */
						uint32_t m = haveoct;
						for( int32_t k = 0; k < nth; k++ ) m &= m - 1;
						sidx = __ffs( (int)m ) - 1;
					}
					o = q - info[sidx].oct0;
					kind = info[sidx].kind;
					floaty = kind == RC_KIND_SOLID || kind == RC_KIND_TURB;
					if( floaty )
					{
						const rc_spanext_t &ex = ext[sidx];
						const bool turbk = kind == RC_KIND_TURB;
						px = turbk ? 16 : 8;
						j = turbk ? o >> 1 : o;
						sdivz = chain[3 * q]; tdivz = chain[3 * q + 1]; zi = chain[3 * q + 2];
						s_raw = ex.s0; t_raw = ex.t0;
						if( j )
							rc_texel_at( ex, sdivz, tdivz, zi, px, &s_raw, &t_raw );
					}
				}
				/* snext / tnext: the neighbour lane holds subdivision j + 1 of the same span unless this is the span's last one */
				int32_t snext = __shfl_down_sync( 0xFFFFFFFFu, s_raw, 1 );
				int32_t tnext = __shfl_down_sync( 0xFFFFFFFFu, t_raw, 1 );
				if( live )
				{
					const rc_spaninfo_t sp = info[sidx];
					rc_octet_t oc;
					const int32_t p0 = o << 3; /* first pixel of the octet inside the span */
					const int32_t npx = min( 8, sp.count - p0 );
					oc.izistep = sp.izistep;
					oc.izi = (int32_t)((uint32_t)sp.izi0 + (uint32_t)p0 * (uint32_t)sp.izistep );
					oc.base = sp.base;
					oc.where = (uint32_t)( sp.x0 + p0 ) | ((uint32_t)npx << 13 ) | ((uint32_t)kind << 17 ) | ((uint32_t)sp.width << 19 );
					oc.s = oc.t = oc.sstep = oc.tstep = 0;
					if( floaty )
					{
						const rc_spanext_t &ex = ext[sidx];
						const bool turbk = kind == RC_KIND_TURB;
						const int32_t spancount = sp.count - j * px;
						const bool last = spancount <= px;
						/* the neighbour's value is the next subdivision's only for 8-pixel subdivisions inside one span */
						const bool own = last ? spancount > 1 : ( turbk || lane == 31 || q + 1 >= total );
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
								const float fpx = (float)px;
								a = sdivz + ex.sdivzstepu * fpx;
								b = tdivz + ex.tdivzstepu * fpx;
								d = zi + ex.zistepu * fpx;
							}
							rc_texel_at( ex, a, b, d, px, &snext, &tnext );
						}
						/* D_DrawSpans16 never reloads s from snext: every later subdivision start is congruent to s0 modulo 8
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
						if( turbk )
						{
							s &= ( RC_CYCLE << 16 ) - 1;
							t &= ( RC_CYCLE << 16 ) - 1;
							if( o & 1 )
							{
								/* second half of a 16-pixel subdivision */
								s = (int32_t)((uint32_t)s + 8u * (uint32_t)sstep );
								t = (int32_t)((uint32_t)t + 8u * (uint32_t)tstep );
							}
						}
						oc.s = s; oc.t = t; oc.sstep = sstep; oc.tstep = tstep;
					}
					uint4 *dst = reinterpret_cast<uint4 *>( &octs[lane] );
					dst[0] = make_uint4( (uint32_t)oc.s, (uint32_t)oc.t, (uint32_t)oc.sstep, (uint32_t)oc.tstep );
					dst[1] = make_uint4( (uint32_t)oc.izi, (uint32_t)oc.izistep, (uint32_t)oc.base, oc.where );
				}
				__syncwarp();

				/* D2: two lanes per octet, four pixels each; 16 octets per pass */
				const int32_t nlive = min( 32, total - r0 );
#pragma unroll
				for( int pass = 0; pass < 2; pass++ )
				{
					const int32_t oi = pass * 16 + ( lane >> 1 );
					if( oi < nlive )
					{
						const uint4 *src = reinterpret_cast<const uint4 *>( &octs[oi] );
						const uint4 A = src[0], B = src[1];
						const int32_t x0 = (int32_t)( B.w & 0x1FFFu ), npx = (int32_t)(( B.w >> 13 ) & 15u );
						const int32_t kind2 = (int32_t)(( B.w >> 17 ) & 3u ), pitch = (int32_t)( B.w >> 19 );
						const int32_t i0 = ( lane & 1 ) << 2;
						uint32_t s = A.x + (uint32_t)i0 * A.z, t = A.y + (uint32_t)i0 * A.w;
						uint32_t izi = B.x + (uint32_t)i0 * B.y;
						const int32_t nmine = min( 4, npx - i0 ); /* pixels of this lane: 4 almost always */
						uint16_t *cp = crow + x0 + i0, *zp = zrow + x0 + i0;
						uint32_t col[4];
						if( kind2 == RC_KIND_SOLID && texfilt )
						{
							/* sw_texfilt 1: dithered texel coordinates (rc_texfilt_index) */
							const uint16_t *blk = cache + (int32_t)B.z;
							const int32_t ln2 = x0 / w.width, X0 = ( x0 - ln2 * w.width ) + npx - i0, Y = ( y + ln2 ) & 1;
#pragma unroll
							for( int i = 0; i < 4; i++ )
							{
								col[i] = i < nmine ? (uint32_t)__ldg( blk + rc_texfilt_index( (int32_t)s, (int32_t)t, pitch, ( X0 - i ) & 1, Y )) : 0u;
								s += A.z; t += A.w;
							}
						}
						else if( kind2 == RC_KIND_SOLID )
						{
							/* one path for full and partial quads (a span's last octet): two paths would both be walked by the
							 * whole warp whenever a pass holds a span end - most passes do */
							const uint16_t *blk = cache + (int32_t)B.z;
#pragma unroll
							for( int i = 0; i < 4; i++ )
							{
								const int32_t ti = ((int32_t)s >> 16 ) + ((int32_t)t >> 16 ) * pitch;
								col[i] = (uint32_t)__ldg( blk + ( i < nmine ? ti : 0 ));
								s += A.z; t += A.w;
							}
						}
						else if( kind2 == RC_KIND_TURB )
						{
							const uint16_t *tex = texels + (int32_t)B.z;
#pragma unroll
							for( int i = 0; i < 4; i++ )
							{
								const int32_t ss = (int32_t)s, tt = (int32_t)t;
								const int32_t sturb = (( ss + turb[( tt >> 16 ) & ( RC_CYCLE - 1 )] ) >> 16 ) & 63;
								const int32_t tturb = (( tt + turb[( ss >> 16 ) & ( RC_CYCLE - 1 )] ) >> 16 ) & 63;
								col[i] = __ldg( tex + ( tturb << 6 ) + sturb );
								s += A.z; t += A.w;
							}
						}
						else
						{
							const uint32_t cv = kind2 == RC_KIND_SKY ? ( B.z & 0xFFFFu ) : 0u;
#pragma unroll
							for( int i = 0; i < 4; i++ ) col[i] = cv;
						}
						uint32_t zv[4];
						if( kind2 == RC_KIND_BACKGROUND )
						{
#pragma unroll
							for( int i = 0; i < 4; i++ ) zv[i] = 0u;
						}
						else
						{
#pragma unroll
							for( int i = 0; i < 4; i++ )
							{
								zv[i] = (uint32_t)( (int32_t)izi >> 16 ) & 0xFFFFu;
								izi += B.y;
							}
						}
						if( nmine == 4 && !(( x0 + i0 ) & 3 ) && !( w.width & 3 ))
						{
							/* an aligned quad: one 8-byte store per plane */
							*reinterpret_cast<uint2 *>( cp ) = make_uint2( col[0] | ( col[1] << 16 ), col[2] | ( col[3] << 16 ));
							*reinterpret_cast<uint2 *>( zp ) = make_uint2( zv[0] | ( zv[1] << 16 ), zv[2] | ( zv[3] << 16 ));
						}
						else if( nmine == 4 && !(( x0 + i0 ) & 1 ) && !( w.width & 1 ))
						{
							/* an even start: two 4-byte stores per plane */
							reinterpret_cast<uint32_t *>( cp )[0] = col[0] | ( col[1] << 16 ); reinterpret_cast<uint32_t *>( cp )[1] = col[2] | ( col[3] << 16 );
							reinterpret_cast<uint32_t *>( zp )[0] = zv[0] | ( zv[1] << 16 ); reinterpret_cast<uint32_t *>( zp )[1] = zv[2] | ( zv[3] << 16 );
						}
						else if( nmine == 4 && !( w.width & 1 ))
						{
							/* an odd start: pixel, aligned pair, pixel */
							cp[0] = (uint16_t)col[0]; *reinterpret_cast<uint32_t *>( cp + 1 ) = col[1] | ( col[2] << 16 ); cp[3] = (uint16_t)col[3];
							zp[0] = (uint16_t)zv[0]; *reinterpret_cast<uint32_t *>( zp + 1 ) = zv[1] | ( zv[2] << 16 ); zp[3] = (uint16_t)zv[3];
						}
						else
						{
#pragma unroll
							for( int i = 0; i < 4; i++ )
								if( i < nmine )
								{
									cp[i] = (uint16_t)col[i];
									zp[i] = (uint16_t)zv[i];
								}
						}
					}
				}
				__syncwarp();
			}
		}
	}
}

#endif /* RC_K_DRAW_CUH */
