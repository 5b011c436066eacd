/*
 * rc_k_scan.cuh - S3, the scanline stage: R_ScanEdges / R_GenerateSpans (ref/soft/r_edge.c:360-675) and the
 * span drawers D_DrawSpans16 / D_DrawZSpans / Turbulent8 / D_BackgroundSurf (r_scan.c:89-221,528-668,1309;
 * r_edge.c:823) as one kernel.
 *
 * One block owns SC_LINES consecutive scanlines of one view and works in two phases per pass:
 *
 *   span phase  (thread per scanline, two threads share a line's edge filtering)
 *     the view's edge list is streamed through shared memory; every line keeps the edges that cross it
 *     as events {u at this line, surfaces} in an exact-fit shared-memory slice (sizes from a block prefix
 *     sum), insertion-sorts them by (u, surfaces) and replays them against the key-ordered surface stack
 *     (rc_generate_spans).  The spans overwrite the line's event slice.
 *   draw phase  (warp per scanline)
 *     lanes set up the 8-pixel (16 for turbulent) perspective chunks of the line's spans into a
 *     shared-memory table (one IEEE divide pair per chunk), then walk the line in aligned groups of 8
 *     pixels: texel gathers from the surface cache, one 16-byte colour store and one 16-byte z store per
 *     lane - fully coalesced 512-byte rows per warp instruction.
 *
 * A pass takes as many of the block's lines as fit the event capacity; normally that is all of them.
 */
#ifndef RC_K_SCAN_CUH
#define RC_K_SCAN_CUH

#define SC_LINES    64
#define SC_THREADS  128
#define SC_WARPS    ( SC_THREADS / 32 )
#define SC_ECHUNK   256  /* edges staged per round */
#define SC_SLOTS    2048 /* event/span slots per pass (each line uses its event count + 1) */
#define SC_SPCAP    128  /* spans per draw batch */
#define SC_BITWORDS 64   /* words of the per-line surface bit set kept in shared memory (<= 2015 faces in view); else global */
#define SC_BUCKETS  128  /* u buckets of the per-line counting sort (uint16 counts share the bit-set storage) */
#define SC_EXCCAP   8

struct rc_spaninfo_t
{
	int32_t u, end;      /* pixel range [u, end) */
	int32_t chunk0;      /* first entry of the span in the chunk table */
	int32_t kind;        /* RC_KIND_* */
	int32_t base, width; /* texel offset + row pitch of the cache block / texture; sky: base = colour */
	int32_t izi0, izistep;
}; /* 32 B */

struct rc_scan_layout_t
{
	int32_t chunkcap;    /* chunk table entries per warp */
	int32_t unionbytes;  /* span-phase scratch (sorted events, bucket counts / surface bit sets) == draw-phase scratch */
	size_t  bytes;
};

static inline rc_scan_layout_t rc_scan_layout( int width )
{
	rc_scan_layout_t l;
	l.chunkcap = ( width + 7 ) / 8 + SC_SPCAP + 1;
	size_t span_u = 8 * (size_t)SC_SLOTS + (size_t)SC_LINES * SC_BITWORDS * 4;
	size_t draw_u = (size_t)SC_WARPS * ( sizeof( rc_spaninfo_t ) * SC_SPCAP + sizeof( int4 ) * (size_t)l.chunkcap );
	l.unionbytes = (int32_t)(( span_u > draw_u ? span_u : draw_u ) + 15 ) & ~15;
	l.bytes = sizeof( rc_edge_t ) * SC_ECHUNK                /* staged edges */
		+ 8 * (size_t)SC_SLOTS                               /* events -> spans */
		+ (size_t)l.unionbytes
		+ sizeof( int32_t ) * ( SC_THREADS + 2 * SC_LINES + 8 ); /* counts, line offsets, span counts */
	return l;
}

__device__ __forceinline__ int32_t rc_warp_incl_scan( int32_t x, int lane )
{
	for( int s = 1; s < 32; s <<= 1 )
	{
		int32_t t = __shfl_up_sync( 0xFFFFFFFFu, x, s );
		if( lane >= s ) x += t;
	}
	return x;
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


__device__ __forceinline__ int32_t rc_key_bucket( unsigned long long key, int32_t x0, int32_t bshift, int32_t nbk )
{
	int32_t iu = ((int32_t)((uint32_t)( key >> 32 ) ^ 0x80000000u )) >> 20;
	int32_t b = ( iu - x0 ) >> bshift;
	return b < 0 ? 0 : ( b >= nbk ? nbk - 1 : b );
}

__global__ void __launch_bounds__( SC_THREADS ) k_scanlines( rc_scene_t sc, rc_work_t w, int32_t chunkcap, int32_t unionbytes )
{
	extern __shared__ __align__( 16 ) unsigned char smem_raw[];
	unsigned char *p = smem_raw;
	rc_edge_t *s_edges = reinterpret_cast<rc_edge_t *>( p ); p += sizeof( rc_edge_t ) * SC_ECHUNK;
	unsigned long long *s_slots = reinterpret_cast<unsigned long long *>( p ); p += 8 * (size_t)SC_SLOTS;
	unsigned char *s_union = p; p += unionbytes;
	int32_t *s_tc = reinterpret_cast<int32_t *>( p ); p += 4 * SC_THREADS;
	int32_t *s_off = reinterpret_cast<int32_t *>( p ); p += 4 * ( SC_LINES + 1 );
	int32_t *s_nsp = reinterpret_cast<int32_t *>( p ); p += 4 * SC_LINES;
	int32_t *s_misc = reinterpret_cast<int32_t *>( p );
	/* span phase view of the union */
	unsigned long long *s_tmp = reinterpret_cast<unsigned long long *>( s_union );
	unsigned char *s_bitbase = s_union + 8 * (size_t)SC_SLOTS;
	/* draw phase view of the union */
	rc_spaninfo_t *s_info = reinterpret_cast<rc_spaninfo_t *>( s_union );
	int4 *s_tab = reinterpret_cast<int4 *>( s_union + sizeof( rc_spaninfo_t ) * SC_SPCAP * SC_WARPS );

	const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
	const int line = t >> 1, half = t & 1;
	const int view = blockIdx.y;
	const int y = blockIdx.x * SC_LINES + line;
	const rc_view_t *vw = &w.views[view];
	const int32_t x0 = vw->vrect[0], y0 = vw->vrect[1], vw_w = vw->vrect[2], vw_h = vw->vrect[3];
	const bool valid = y >= y0 && y < y0 + vw_h && y < w.height;
	int32_t ne = w.nedges[view];
	if( ne > w.max_edges ) ne = w.max_edges;
	const rc_edge_t *gedges = w.edges + (size_t)view * w.max_edges;
	const rc_surf_t *vsurfs = w.surfs + (size_t)view * w.max_cands;
	if( t == 0 ) { s_misc[0] = 0; s_misc[1] = 0; }

	/* surface bit sets: shared memory when the view's face list is short enough, else this block's global scratch */
	const int32_t nwords = ( w.ncands[view] + 32 ) >> 5;
	uint32_t *bitsbase = reinterpret_cast<uint32_t *>( s_bitbase );
	int32_t bitstride = SC_BITWORDS;
	if( nwords > SC_BITWORDS )
	{
		bitstride = w.scanbits_words;
		bitsbase = w.scanbits + ((size_t)view * gridDim.x + blockIdx.x ) * SC_LINES * (size_t)bitstride;
	}
	int32_t bshift = 3;
	while((( vw_w + 1 ) >> bshift ) + 1 > SC_BUCKETS ) bshift++;
	const int32_t nbk = (( vw_w + 1 ) >> bshift ) + 1;

	/* pass 1: how many edges cross each line */
	int32_t cnt = 0;
	for( int32_t base = 0; base < ne; base += SC_ECHUNK )
	{
		const int32_t m = min( SC_ECHUNK, ne - base );
		for( int32_t k = t; k < m; k += SC_THREADS )
			reinterpret_cast<int4 *>( s_edges )[k] = reinterpret_cast<const int4 *>( gedges + base )[k];
		__syncthreads();
		if( valid )
			for( int32_t k = half; k < m; k += 2 )
			{
				const uint32_t vv = s_edges[k].vv;
				cnt += ( y >= (int32_t)( vv & 0xFFFF ) && y <= (int32_t)( vv >> 16 )) ? 1 : 0;
			}
		__syncthreads();
	}
	s_tc[t] = cnt;
	__syncthreads();
	if( t == 0 )
	{
		int32_t off = 0;
		for( int l = 0; l < SC_LINES; l++ )
		{
			s_off[l] = off;
			off += s_tc[2 * l] + s_tc[2 * l + 1] + 1;
		}
		s_off[SC_LINES] = off;
	}
	__syncthreads();

	int32_t l0 = 0;
	while( l0 < SC_LINES )
	{
		const int32_t basev = s_off[l0];
		int32_t l1 = l0 + 1;
		while( l1 < SC_LINES && s_off[l1 + 1] - basev <= SC_SLOTS )
			l1++;
		const bool inpass = valid && line >= l0 && line < l1;
		const int32_t lo = s_off[line] - basev;              /* first slot of this line */
		const int32_t nline = min( s_tc[2 * line] + s_tc[2 * line + 1], SC_SLOTS - 1 ); /* a single line over capacity is truncated */
		int32_t wr = lo + ( half ? s_tc[2 * line] : 0 );
		const int32_t wend = lo + nline;

		/* pass 2: the events of the lines of this pass */
		for( int32_t base = 0; base < ne; base += SC_ECHUNK )
		{
			const int32_t m = min( SC_ECHUNK, ne - base );
			for( int32_t k = t; k < m; k += SC_THREADS )
				reinterpret_cast<int4 *>( s_edges )[k] = reinterpret_cast<const int4 *>( gedges + base )[k];
			__syncthreads();
			if( inpass )
				for( int32_t k = half; k < m; k += 2 )
				{
					const rc_edge_t e = s_edges[k];
					if( y >= (int32_t)( e.vv & 0xFFFF ) && y <= (int32_t)( e.vv >> 16 ))
					{
						if( wr < wend )
							s_slots[wr] = rc_event_key( rc_edge_u_at( &e, y ), e.surfs );
						wr++;
					}
				}
			__syncthreads();
		}

		/* sort by (u, surfaces): counting sort into u buckets, then insertion sort of the nearly sorted result */
		if( inpass && half == 0 && nline > 0 )
		{
			const unsigned long long *a = s_slots + lo;
			unsigned long long *b = s_tmp + lo;
			uint16_t *bk = reinterpret_cast<uint16_t *>( s_bitbase ) + line * SC_BUCKETS;
			for( int32_t k = 0; k < nbk; k++ ) bk[k] = 0;
			for( int32_t i = 0; i < nline; i++ ) bk[rc_key_bucket( a[i], x0, bshift, nbk )]++;
			int32_t run = 0;
			for( int32_t k = 0; k < nbk; k++ ) { const int32_t c = bk[k]; bk[k] = (uint16_t)run; run += c; }
			for( int32_t i = 0; i < nline; i++ )
			{
				const unsigned long long key = a[i];
				b[bk[rc_key_bucket( key, x0, bshift, nbk )]++] = key;
			}
			for( int32_t i = 1; i < nline; i++ )
			{
				const unsigned long long key = b[i];
				int32_t j = i - 1;
				while( j >= 0 && b[j] > key ) { b[j + 1] = b[j]; j--; }
				b[j + 1] = key;
			}
		}
		__syncthreads();
		/* the bucket counters become the surface bit sets */
		for( int32_t l = l0 + warp; l < l1; l += SC_WARPS )
			for( int32_t k = lane; k < nwords; k += 32 )
				bitsbase[(size_t)l * bitstride + k] = 0;
		__syncthreads();

		/* replay the sorted edges against the surface set; spans go to the line's slot slice */
		if( half == 0 && line >= l0 && line < l1 )
		{
			int32_t nsp = 0;
			if( valid )
			{
				rc_event_t *ev = reinterpret_cast<rc_event_t *>( s_tmp + lo );
				for( int32_t i = 0; i < nline; i++ )
				{
					const unsigned long long k = s_tmp[lo + i];
					rc_event_t e;
					e.u = (int32_t)((uint32_t)( k >> 32 ) ^ 0x80000000u );
					e.surfs = (uint32_t)k;
					ev[i] = e;
				}
				int32_t exc[2 * SC_EXCCAP];
				int32_t ovf = ( s_tc[2 * line] + s_tc[2 * line + 1] > nline ) ? 1 : 0;
				nsp = rc_generate_spans_bits( ev, nline, x0, (( x0 + vw_w ) << 20 ) + 0xFFFFF, bitsbase + (size_t)line * bitstride, nwords,
					exc, SC_EXCCAP, reinterpret_cast<rc_span_t *>( s_slots + lo ), nline + 1, &ovf );
				atomicAdd( &s_misc[0], nsp );
				if( ovf ) atomicAdd( &s_misc[1], 1 );
			}
			s_nsp[line] = nsp;
		}
		__syncthreads();

		/* draw: warp per line (the scratch of the span phase is dead now) */
		for( int32_t l = l0 + warp; l < l1; l += SC_WARPS )
		{
			const int32_t nsp = s_nsp[l];
			if( nsp <= 0 )
				continue;
			const int32_t yy = blockIdx.x * SC_LINES + l;
			rc_draw_line( sc, w, vw, vsurfs, reinterpret_cast<const rc_span_t *>( s_slots + ( s_off[l] - basev )), nsp, yy,
				s_info + warp * SC_SPCAP, s_tab + (size_t)warp * chunkcap,
				w.color + ((size_t)view * w.height + yy ) * w.width, w.zbuf + ((size_t)view * w.height + yy ) * w.width, lane );
		}
		__syncthreads();
		l0 = l1;
	}

	if( t == 0 )
	{
		if( s_misc[0] ) atomicAdd( &w.counters[RC_CTR_SPANS], (unsigned long long)s_misc[0] );
		if( s_misc[1] ) atomicAdd( &w.counters[RC_CTR_OVERFLOW_EVENTS], (unsigned long long)s_misc[1] );
	}
}

#endif /* RC_K_SCAN_CUH */
