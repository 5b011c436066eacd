/*
 * rc_k_spans.cuh - S3a, span generation: R_ScanEdges / R_InsertNewEdges / R_StepActiveU / R_GenerateSpans /
 * R_LeadingEdge / R_TrailingEdge / R_CleanupSpan (ref/soft/r_edge.c:122-675) for every scanline of every view.
 *
 * The reference sweeps a view top to bottom keeping one sorted active-edge list.  An edge's u at scanline y is
 * closed form (u0 + (y - v) * u_step in int32 wraparound, r_edge.c:187), so scanlines are independent given
 * the edge set.  Two kernels:
 *
 *   k_events  block per (view, band of SP_LINES = 32 scanlines), bands issued from the middle of the frame outwards:
 *             gathers the view's edges that cross the band (ballot compaction), pre-sorts them by u at the band's
 *             middle line, renumbers the band's surfaces locally, permutes the edges into sorted order; then lane =
 *             scanline, warps take runs of the sorted list: pass 1 counts the edges crossing every line, a prefix sum
 *             gives every line its slice of the band's arena in HBM, pass 2 writes the event keys {u at the line,
 *             band-local surfaces} - nearly sorted already.  Lines with more than LN_HEAVY events go to a list.
 *   k_lines   warp per scanline, lines drawn from shared cursors (the heavy list first): sortedness check, odd-even
 *             passes / bitonic network (registers + shuffles up to 32 events, shared memory up to SP_SORTCAP) = the
 *             order R_InsertNewEdges / R_StepActiveU keep incrementally, then the replay against the surface bit set
 *             by the whole warp (rc_replay_warp: XOR scan, 32 events per round; rc_generate_spans_bits for degenerate
 *             lines); the spans overwrite the slice, the count goes to the line table k_draw reads.
 */
#ifndef RC_K_SPANS_CUH
#define RC_K_SPANS_CUH

/* SP_LINES (scanlines per band) is defined by rc_device.cu */
/* k_events is built for 256 threads (frames up to 640 wide: a few hundred edges per band) and for 512 / 1024 (wide frames of
 * large maps: thousands of edges per band, few blocks - C3: spans 1.38 -> 1.01 ms with 512; 256 stays faster on C2 / C5) */
#define LN_WARPS    4
#define LN_HEAVY    128  /* events: lines above are listed by k_events and go first (measured: 32: 166 us, 64: 159, 128: 153, 192: 159) */
#define LN_CHUNK    1    /* scanlines a warp takes from the shared cursor at a time (measured on the C2 batch: 1: 170 us, 2: 172, 4: 183, 8: 184) */
#define SP_SORTCAP  512  /* events per line k_lines sorts in shared memory; longer lines go to k_lines_wide */
#define WD_SORTCAP  4096 /* events per line k_lines_wide sorts in shared memory (a block per line); beyond: serial */
#define WD_THREADS  128
#define WD_GRID     ( 148 * 6 )
#define RC_LINE_WIDE 0x20000000 /* set in nsp by k_events: more events than k_lines sorts or more band-local surfaces than its
                                 * register bit sets hold - the line is k_lines_wide's */
#define SP_BITWORDS 64   /* shared-memory bit-set words per line (views listing <= 2015 faces), else global scratch */
#define SP_EXCCAP   8
#define SP_OEPASSES 6  /* odd-even transposition passes tried on a nearly sorted line before the bitonic network */

#if SP_LINES != 32
#error "k_events maps lanes to scanlines: SP_LINES must be 32"
#endif

/* band edges staged and pre-sorted in shared memory: w.ev_ecap of them (24 B each: the edge and its sort key); longer
 * lists are read from the view's edge list in HBM, unsorted.  The presort key has 13 bits for the edge number. */
#define EV_ECAP_SMALL 2048
#define EV_ECAP_BIG   4096

static inline size_t rc_events_smem( int32_t ecap, int32_t bitwords )
{
	return sizeof( rc_edge_t ) * (size_t)ecap + sizeof( unsigned long long ) * (size_t)ecap + sizeof( uint32_t ) * 2 * ( (size_t)bitwords + 1 );
}

template <int EV_THREADS>
__global__ void __launch_bounds__( EV_THREADS, 1024 / EV_THREADS ) k_events( rc_work_t w )
{
	extern __shared__ __align__( 16 ) unsigned char ev_smem[];
	__shared__ int32_t s_wcnt[( EV_THREADS / 32 )][SP_LINES];
	__shared__ int32_t s_warp[( EV_THREADS / 32 )];
	__shared__ int32_t s_ne;
	rc_edge_t *s_edge = reinterpret_cast<rc_edge_t *>( ev_smem );
	unsigned long long *s_key = reinterpret_cast<unsigned long long *>( ev_smem + sizeof( rc_edge_t ) * (size_t)w.ev_ecap );
	uint32_t *s_bm = reinterpret_cast<uint32_t *>( ev_smem + ( sizeof( rc_edge_t ) + sizeof( unsigned long long )) * (size_t)w.ev_ecap );

	const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
	/* bands are issued from the middle of the frame outwards: the horizon of an indoor view is where the edges are
	 * (bands with 600 edges next to bands with 30), and the longest blocks should start first */
	const int view = blockIdx.x, nbands = gridDim.y;
	const int band = ( nbands >> 1 ) + (( blockIdx.y & 1 ) ? -(int)(( blockIdx.y + 1 ) >> 1 ) : (int)( blockIdx.y >> 1 ));
#ifdef RC_SPANS_DEBUG
	long long ev_t[10]; int ev_n = 0;
#define EV_MARK() do { __syncthreads(); ev_t[ev_n++] = clock64(); } while( 0 )
#else
#define EV_MARK() do { } while( 0 )
#endif
	EV_MARK();
	const rc_view_t *vw = &w.views[view];
	const int32_t y0 = vw->vrect[1], vw_h = vw->vrect[3];
	const int32_t y = band * SP_LINES + lane;
	const bool valid = y >= y0 && y < y0 + vw_h && y < w.height;

	const int32_t nview = min( w.nedges[view], w.max_edges );
	if( t == 0 && blockIdx.y == 0 ) atomicMax( &w.counters[RC_CTR_MAX_EDGES], (unsigned long long)w.nedges[view] );
	const rc_edge_t *vedges = w.edges + (size_t)view * w.max_edges;
	unsigned long long *arena = w.events + ((size_t)view * nbands + band ) * w.blk_events;
	const int32_t bmwords = ( w.ncands[view] + 32 ) >> 5; /* surface numbers are candidate index + 1 */
	uint32_t *s_pre = s_bm + bmwords;

	/* A. gather the view's edges that cross this band (a view has a few hundred edges: every band block reads them
	 * all, which is cheaper than k_faces scattering them into band lists with atomics), mark the surfaces they belong
	 * to, sort keys = u at the band's middle line (taken at the nearest line the edge is active on) */
	const int32_t yb0 = band * SP_LINES, yb1 = yb0 + SP_LINES - 1;
	if( t == 0 ) s_ne = 0;
	for( int32_t i = t; i < bmwords; i += EV_THREADS ) s_bm[i] = 0;
	__syncthreads();
	for( int32_t base = 0; base < nview; base += EV_THREADS )
	{
		const int32_t k = base + t;
		rc_edge_t e;
		bool in = false;
		if( k < nview )
		{
			e = vedges[k];
			const int32_t v = (int32_t)( e.vv & 0xFFFF ), v2 = (int32_t)( e.vv >> 16 );
			in = v <= v2 && v <= yb1 && v2 >= yb0; /* reservations a face did not use have v > v2 */
		}
		const uint32_t m = __ballot_sync( 0xFFFFFFFFu, in );
		int32_t wb = 0;
		if( lane == 0 && m ) wb = atomicAdd( &s_ne, __popc( m ));
		wb = __shfl_sync( 0xFFFFFFFFu, wb, 0 );
		if( in )
		{
			const int32_t slot = wb + __popc( m & (( 1u << lane ) - 1u ));
			if( slot < w.ev_ecap ) s_edge[slot] = e;
			const int32_t s = (int32_t)(( e.surfs & 0xFFFF ) ? ( e.surfs & 0xFFFF ) : ( e.surfs >> 16 ));
			atomicOr( &s_bm[s >> 5], 1u << ( s & 31 ));
		}
	}
	__syncthreads();
	EV_MARK(); /* 1: gathered */
	/* more edges than the staging area holds: the lines read the view's edge list from HBM, unsorted */
	const bool staged = s_ne <= w.ev_ecap;
	if( t == 0 && !staged ) atomicAdd( &w.counters[RC_CTR_UNSTAGED_BANDS], 1ull );
	const int32_t ne = staged ? s_ne : nview;
	int32_t npow = 32;
	while( npow < ne ) npow <<= 1;
	const int32_t ymid = band * SP_LINES + SP_LINES / 2;
	if( staged )
		for( int32_t k = t; k < npow; k += EV_THREADS )
		{
			if( k >= ne ) { s_key[k] = ~0ull; continue; }
			const rc_edge_t e = s_edge[k];
			const int32_t s = (int32_t)(( e.surfs & 0xFFFF ) ? ( e.surfs & 0xFFFF ) : ( e.surfs >> 16 ));
			const int32_t v = (int32_t)( e.vv & 0xFFFF ), v2 = (int32_t)( e.vv >> 16 );
			const int32_t ym = ymid < v ? v : ( ymid > v2 ? v2 : ymid );
			const uint32_t um = (uint32_t)e.u + (uint32_t)( ym - v ) * (uint32_t)e.u_step;
			/* ties (the two copies of a shared edge) in the order the event keys will have: leading first, then by surface */
			s_key[k] = ((unsigned long long)( um ^ 0x80000000u ) << 32 ) | (( e.surfs & 0xFFFF ) ? 1u << 29 : 0u ) | ((uint32_t)s << 13 ) | (uint32_t)k;
		}
	__syncthreads();

	EV_MARK(); /* 2: keys */
	/* B. block bitonic sort of the keys: the order the lines will (nearly) need */
	if( staged && ne > 1 && ne <= EV_THREADS )
	{
		/* short lists (the usual case): rank sort - every thread counts the keys below its own (broadcast reads,
		 * no barriers inside), keys are unique (the edge number is part of them) */
		const unsigned long long mine = t < ne ? s_key[t] : 0ull;
		int32_t rank = 0;
		if( t < ne )
			for( int32_t i = 0; i < ne; i++ ) rank += s_key[i] < mine ? 1 : 0;
		__syncthreads();
		if( t < ne ) s_key[rank] = mine;
		__syncthreads();
	}
	else if( staged && ne > 1 )
		for( int32_t k = 2; k <= npow; k <<= 1 )
			for( int32_t j = k >> 1; j > 0; j >>= 1 )
			{
				for( int32_t pr = t; pr < ( npow >> 1 ); pr += EV_THREADS )
				{
					const int32_t a = (( pr & ~( j - 1 )) << 1 ) | ( pr & ( j - 1 )), b = a + j;
					const unsigned long long va = s_key[a], vb = s_key[b];
					if(( va > vb ) == (( a & k ) == 0 )) { s_key[a] = vb; s_key[b] = va; }
				}
				__syncthreads();
			}

	EV_MARK(); /* 3: sorted */
	/* C. band-local surface numbers: rank of the surface among those present in the band, + 1 (order preserving, so
	 * "lowest set bit = smallest key" still holds in k_lines, with far fewer bit-set words) */
	int32_t carry = 0;
	for( int32_t i0 = 0; i0 < bmwords; i0 += EV_THREADS )
	{
		const int32_t i = i0 + t;
		const int32_t v = i < bmwords ? __popc( s_bm[i] ) : 0;
		const int32_t x = rc_warp_incl_scan( v, lane );
		if( lane == 31 ) s_warp[warp] = x;
		__syncthreads();
		int32_t wbase = 0, tot = 0;
		for( int k = 0; k < ( EV_THREADS / 32 ); k++ )
		{
			if( k < warp ) wbase += s_warp[k];
			tot += s_warp[k];
		}
		if( i < bmwords ) s_pre[i] = (uint32_t)( carry + wbase + x - v );
		carry += tot;
		__syncthreads();
	}
	const int32_t nlocal = carry;
	uint16_t *bsurf = w.bandsurf + ((size_t)view * nbands + band ) * ( (size_t)w.max_cands + 2 );
	for( int32_t i = t; i < bmwords; i += EV_THREADS )
	{
		uint32_t m = s_bm[i];
		int32_t r = (int32_t)s_pre[i] + 1;
		while( m )
		{
			bsurf[r++] = (uint16_t)(( i << 5 ) + __ffs( (int)m ) - 1 );
			m &= m - 1;
		}
	}
	if( t == 0 )
	{
		bsurf[0] = 0;
		w.bandnloc[view * nbands + band] = nlocal;
	}
	/* the staged edges are permuted into sorted order (through registers: all reads, barrier, all writes) with the
	 * band-local surface number in the lead / trail half of surfs: the per-line loops below then read one 16-byte
	 * record per edge and nothing else */
	if( staged )
	{
		rc_edge_t mine[EV_ECAP_BIG / EV_THREADS];
#pragma unroll
		for( int q = 0; q < EV_ECAP_BIG / EV_THREADS; q++ )
		{
			const int32_t k = q * EV_THREADS + t;
			if( k < ne )
			{
				rc_edge_t e = s_edge[(int32_t)( (uint32_t)s_key[k] & 8191u )];
				const uint32_t tr = e.surfs & 0xFFFF, s = tr ? tr : ( e.surfs >> 16 );
				const uint32_t loc = s_pre[s >> 5] + (uint32_t)__popc( s_bm[s >> 5] & (( 1u << ( s & 31 )) - 1u )) + 1u;
				e.surfs = tr ? loc : ( loc << 16 );
				mine[q] = e;
			}
		}
		__syncthreads();
#pragma unroll
		for( int q = 0; q < EV_ECAP_BIG / EV_THREADS; q++ )
		{
			const int32_t k = q * EV_THREADS + t;
			if( k < ne ) s_edge[k] = mine[q];
		}
		__syncthreads();
	}
	const rc_edge_t *L = staged ? s_edge : vedges;

	EV_MARK(); /* 4: local numbers */
	/* 1. events per line: every warp takes a contiguous run of the sorted edge list */
	const int32_t per = ( ne + ( EV_THREADS / 32 ) - 1 ) / ( EV_THREADS / 32 );
	const int32_t k0 = warp * per, k1 = min( ne, k0 + per );
	int32_t cnt = 0;
	for( int32_t k = k0; k < k1; k++ )
	{
		const uint32_t vv = L[k].vv;
		cnt += ( valid && y >= (int32_t)( vv & 0xFFFF ) && y <= (int32_t)( vv >> 16 )) ? 1 : 0;
	}
	s_wcnt[warp][lane] = cnt;
	__syncthreads();
	int32_t mybase = 0, n = 0;
	for( int k = 0; k < ( EV_THREADS / 32 ); k++ )
	{
		const int32_t c = s_wcnt[k][lane];
		if( k < warp ) mybase += c;
		n += c;
	}
	/* every line gets its event count + 1 slots (the spans that replace the events can be one more) */
	const int32_t incl = rc_warp_incl_scan( n + 1, lane );
	const int32_t off = incl - ( n + 1 );
	const bool fits = off + n + 1 <= w.blk_events;
	if( warp == 0 )
	{
		if( lane == 31 ) atomicMax( &w.counters[RC_CTR_MAX_BAND_EVENTS], (unsigned long long)incl );
		rc_lineinfo_t li;
		li.ofs = -1; li.nsp = 0;
		if( valid && fits )
		{
			li.ofs = band * w.blk_events + off; /* index inside the view's arena */
			li.nsp = n;                         /* events for now; k_lines replaces it with the span count */
			if( n > SP_SORTCAP || (( nlocal + 32 ) >> 5 ) > 16 )
			{
				li.nsp |= RC_LINE_WIDE;
				w.wide_list[atomicAdd( w.wide_count, 1 )] = view * w.height + y;
			}
			else if( n > LN_HEAVY ) w.heavy_list[atomicAdd( w.heavy_count, 1 )] = view * w.height + y;
		}
		else if( valid )
			atomicAdd( &w.counters[RC_CTR_OVERFLOW_EVENTS], 1ull );
		if( y < w.height )
			w.lineinfo[(size_t)view * w.height + y] = li;
	}
	EV_MARK(); /* 5: counted */
#ifndef RC_SPANS_DEBUG
	if( !valid || !fits )
		return;
#endif

	/* 2. the events, surfaces renumbered band-locally */
	int32_t pos = off + mybase;
#ifdef RC_SPANS_DEBUG
	if( valid && fits )
#endif
	for( int32_t k = k0; k < k1; k++ )
	{
		const rc_edge_t e = L[k];
		const int32_t v = (int32_t)( e.vv & 0xFFFF );
		if( y >= v && y <= (int32_t)( e.vv >> 16 ))
		{
			uint32_t both = e.surfs;
			if( !staged )
			{
				const uint32_t tr = e.surfs & 0xFFFF, s = tr ? tr : ( e.surfs >> 16 );
				const uint32_t loc = s_pre[s >> 5] + (uint32_t)__popc( s_bm[s >> 5] & (( 1u << ( s & 31 )) - 1u )) + 1u;
				both = tr ? loc : ( loc << 16 );
			}
			arena[pos++] = rc_event_key( (int32_t)((uint32_t)e.u + (uint32_t)( y - v ) * (uint32_t)e.u_step ), both );
		}
	}
	EV_MARK(); /* 6: filled */
#ifdef RC_SPANS_DEBUG
	if( t == 0 )
		for( int k = 0; k < 6; k++ )
		{
			atomicAdd( &w.counters[48 + k], (unsigned long long)( ev_t[k + 1] - ev_t[k] ));
			atomicMax( &w.counters[56 + k], (unsigned long long)( ev_t[k + 1] - ev_t[k] ));
		}
#endif
}

/* bitonic sort of n <= SP_SORTCAP keys in shared memory by one warp; buf holds the keys padded to a power of two */
__device__ __forceinline__ void rc_sort_smem( unsigned long long *buf, int32_t npow, int lane )
{
	for( int32_t k = 2; k <= npow; k <<= 1 )
		for( int32_t j = k >> 1; j > 0; j >>= 1 )
		{
			for( int32_t pr = lane; pr < ( npow >> 1 ); pr += 32 )
			{
				const int32_t a = (( pr & ~( j - 1 )) << 1 ) | ( pr & ( j - 1 )), b = a + j; /* j is a power of two */
				const unsigned long long va = buf[a], vb = buf[b];
				if(( va > vb ) == (( a & k ) == 0 )) { buf[a] = vb; buf[b] = va; }
			}
			__syncwarp();
		}
}

/*
 * R_GenerateSpans for one scanline by a whole warp, lane per event (32 sorted events per round).
 *
 * What the serial replay (rc_generate_spans_bits) carries from event to event is the set of surfaces on the
 * stack; with spanstate in {0, 1} a leading edge sets its surface's bit and a trailing edge clears it, i.e. both
 * TOGGLE it, so the set after event i is the XOR of the toggles of events 0..i: one inclusive XOR scan over the
 * lanes (NW words of bit set per lane, 5 shuffle steps each).  From the set after every event: the top surface
 * (lowest set bit = smallest key), the top before it (neighbour lane), whether the event ends a span (trailing
 * edge of the top / leading edge in front of the top), the start of that span (running max of the pixel column
 * of earlier span ends) - and the spans are compacted out with a ballot.  A leading edge that finds its bit set,
 * or a trailing edge that finds it clear (spanstate outside {0, 1}: inverted or doubled spans), makes the warp
 * give the line to the serial replay instead.  Returns the span count, or -1 for "replay serially".
 */
template <int NW>
__device__ __forceinline__ int32_t rc_replay_warp( const unsigned long long *keys, int32_t n, int32_t head_iu, int32_t tail_u,
	const uint16_t *bsurf, rc_span_t *spans, int lane )
{
	uint32_t st[NW];
#pragma unroll
	for( int k = 0; k < NW; k++ ) st[k] = 0;
	int32_t top_carry = 0, last_u = head_iu, nspans = 0;
	const uint32_t lt = ( 1u << lane ) - 1u;

	for( int32_t c0 = 0; c0 < n; c0 += 32 )
	{
		const int32_t i = c0 + lane;
		const unsigned long long key = i < n ? keys[i] : 0ull;
		const int32_t u = (int32_t)((uint32_t)( key >> 32 ) ^ 0x80000000u );
		const uint32_t both = (uint32_t)key;
		const bool valid = i < n && u <= tail_u;
		const int32_t iu = u >> 20;
		const uint32_t trail = both >> 16, lead = both & 0xFFFFu;
		const bool is_lead = lead != 0;
		const int32_t s = (int32_t)( is_lead ? lead : trail );
		const uint32_t bit = 1u << ( s & 31 );
		const int32_t wsel = s >> 5;
		bool exc = valid && (( trail != 0 && lead != 0 ) || wsel >= NW );

		uint32_t x[NW];
#pragma unroll
		for( int k = 0; k < NW; k++ ) x[k] = ( valid && wsel == k ) ? bit : 0u;
#pragma unroll
		for( int d = 1; d < 32; d <<= 1 )
		{
#pragma unroll
			for( int k = 0; k < NW; k++ )
			{
				const uint32_t o = __shfl_up_sync( 0xFFFFFFFFu, x[k], d );
				if( lane >= d ) x[k] ^= o;
			}
		}
		uint32_t mine = 0;
		int32_t top_after = 0;
#pragma unroll
		for( int k = NW - 1; k >= 0; k-- )
		{
			x[k] ^= st[k]; /* the set after this lane's event */
			if( wsel == k ) mine = x[k];
			if( x[k] ) top_after = ( k << 5 ) + __ffs( (int)x[k] ) - 1;
		}
		/* leading edge: the bit must have been clear (now set); trailing edge: set (now clear) */
		exc = exc || ( valid && ((( mine & bit ) != 0 ) != is_lead ));
		if( __ballot_sync( 0xFFFFFFFFu, exc ))
			return -1;

		int32_t top_before = __shfl_up_sync( 0xFFFFFFFFu, top_after, 1 );
		if( lane == 0 ) top_before = top_carry;
		const bool brk = valid && ( is_lead ? ( top_before == 0 || s < top_before ) : ( s == top_before ));

		/* pixel column of the latest span end before this event */
		int32_t m = brk ? iu : (int32_t)0x80000000;
#pragma unroll
		for( int d = 1; d < 32; d <<= 1 )
		{
			const int32_t o = __shfl_up_sync( 0xFFFFFFFFu, m, d );
			if( lane >= d && o > m ) m = o;
		}
		int32_t prev = __shfl_up_sync( 0xFFFFFFFFu, m, 1 );
		if( lane == 0 || prev < last_u ) prev = last_u;
		const bool emit = brk && iu > prev;
		const uint32_t em = __ballot_sync( 0xFFFFFFFFu, emit );
		if( emit )
		{
			rc_span_t sp;
			sp.surf = bsurf[top_before]; sp.u = (int16_t)prev; sp.count = (int16_t)( iu - prev );
			spans[nspans + __popc( em & lt )] = sp;
		}
		nspans += __popc( em );
#pragma unroll
		for( int k = 0; k < NW; k++ ) st[k] = __shfl_sync( 0xFFFFFFFFu, x[k], 31 );
		top_carry = __shfl_sync( 0xFFFFFFFFu, top_after, 31 );
		const int32_t mlast = __shfl_sync( 0xFFFFFFFFu, m, 31 );
		if( mlast > last_u ) last_u = mlast;
	}
	/* R_CleanupSpan: whatever is on top runs to the right edge */
	const int32_t iu = tail_u >> 20;
	if( iu > last_u )
	{
		if( lane == 0 )
		{
			rc_span_t sp;
			sp.surf = bsurf[top_carry]; sp.u = (int16_t)last_u; sp.count = (int16_t)( iu - last_u );
			spans[nspans] = sp;
		}
		nspans++;
	}
	return nspans;
}

/* brush entities: a scanline that carries a piece of one is replayed by the serial generator that keeps R_LeadingEdge's
 * 1/z order among pieces with one leaf key (rc_generate_spans); views without brush entities never get here */
#define SP_EXACT_STK 96
__device__ __forceinline__ bool rc_line_has_bmodel( const rc_work_t &w, int view, const unsigned long long *keys, int32_t n, const uint16_t *bsurf, int lane, int nlanes )
{
	if( !w.nbents || !w.viewbm[view] )
		return false;
	const rc_surf_t *vs = w.surfs + (size_t)view * w.max_cands;
	bool any = false;
	for( int32_t i = lane; i < n; i += nlanes )
	{
		const uint32_t lead = (uint32_t)keys[i] & 0xFFFFu;
		if( lead ) any = any || ( vs[bsurf[lead] - 1].emitted >> 8 ) != 0;
	}
	return any;
}

__device__ __noinline__ int32_t rc_replay_exact( const rc_work_t &w, int view, int32_t y, const unsigned long long *keys, int32_t n, int32_t head_iu, int32_t tail_u,
	const uint16_t *bsurf, rc_span_t *spans, int32_t *ovf )
{
	int32_t stk[SP_EXACT_STK], cnt[SP_EXACT_STK], neg[2 * SP_EXCCAP];
	return rc_generate_spans( keys, n, head_iu, tail_u, bsurf, stk, cnt, SP_EXACT_STK, neg, SP_EXCCAP, spans, n + 1, ovf,
		w.surfs + (size_t)view * w.max_cands, y );
}

__global__ void __launch_bounds__( LN_WARPS * 32, 8 ) k_lines( rc_work_t w )
{
	__shared__ __align__( 16 ) unsigned long long s_keys[LN_WARPS][SP_SORTCAP];
	__shared__ uint32_t s_bits[LN_WARPS][SP_BITWORDS + 1];

	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const int32_t gwarp = blockIdx.x * LN_WARPS + warp;
	int32_t cur = 0;
	const int32_t nlines = w.nviews * w.height;
	unsigned long long *keys = s_keys[warp];
	int32_t nspans_total = 0, novf = 0, nserial = 0, nsorted = 0, nties = 0;
#ifdef RC_SPANS_DEBUG
	const long long dbg_start = clock64();
#endif

	/* scanlines differ a lot in cost (2 .. 500 events) and the heavy ones cluster (the horizon of corridor views):
	 * warps draw them from a shared cursor, view-interleaved */
	/* two sweeps over the lines: the heavy ones first (longest processing time first keeps the tail short) */
	const int32_t nheavy = *w.heavy_count; /* k_events listed the lines with more than LN_HEAVY events */
	for( int32_t heavy = 1; ; )
	{
		int32_t item = 0;
		if( lane == 0 ) item = atomicAdd( w.line_cursor + ( heavy ? 0 : 1 ), 1 );
		item = __shfl_sync( 0xFFFFFFFFu, item, 0 );
		if( item >= ( heavy ? nheavy : nlines ))
		{
			if( !heavy )
				break;
			heavy = 0;
			continue;
		}
		int32_t view, line;
		if( heavy )
		{
			line = w.heavy_list[item];
			view = line / w.height;
		}
		else
		{
			view = item % w.nviews;
			line = view * w.height + item / w.nviews;
		}
		const rc_lineinfo_t li = w.lineinfo[line];
		if( li.ofs < 0 )
			continue;
		const int32_t n = li.nsp;
		if(( n & ( RC_LINE_DONE | RC_LINE_WIDE )) || ( !heavy && n > LN_HEAVY ))
			continue; /* done in the first sweep, or k_lines_wide's */
		const rc_view_t *vw = &w.views[view];
		const int32_t x0 = vw->vrect[0], vw_w = vw->vrect[2];
#ifdef RC_SPANS_DEBUG
		const long long dbg0 = clock64();
#endif
		unsigned long long *slice = w.events + (size_t)view * w.view_events + li.ofs;
		const unsigned long long *sorted = keys;

		if( n <= 32 )
		{
			unsigned long long v = lane < n ? slice[lane] : ~0ull;
			/* k_events wrote the keys in the band's pre-sorted edge order: most lines are in order already */
			const unsigned long long nx = __shfl_down_sync( 0xFFFFFFFFu, v, 1 );
			unsigned uns = __ballot_sync( 0xFFFFFFFFu, lane < 31 && v > nx );
			if( uns )
			{
				/* nearly sorted (the band's edge order): a few odd-even transposition passes, else the full network */
				nsorted++;
				for( int pass = 0; uns && pass < SP_OEPASSES; pass++ )
				{
					unsigned long long o = __shfl_xor_sync( 0xFFFFFFFFu, v, 1 );
					v = ( lane & 1 ) ? ( v > o ? v : o ) : ( v < o ? v : o );
					const unsigned long long up = __shfl_down_sync( 0xFFFFFFFFu, v, 1 ), dn = __shfl_up_sync( 0xFFFFFFFFu, v, 1 );
					if( lane & 1 ) { if( lane < 31 && up < v ) v = up; }
					else if( lane > 0 && dn > v ) v = dn;
					const unsigned long long nx2 = __shfl_down_sync( 0xFFFFFFFFu, v, 1 );
					uns = __ballot_sync( 0xFFFFFFFFu, lane < 31 && v > nx2 );
				}
				if( uns )
					for( int k = 2; k <= 32; k <<= 1 )
						for( int j = k >> 1; j > 0; j >>= 1 )
						{
							const unsigned long long o = __shfl_xor_sync( 0xFFFFFFFFu, v, j );
							const bool keepmin = (( lane & k ) == 0 ) == (( lane & j ) == 0 );
							v = keepmin ? ( v < o ? v : o ) : ( v > o ? v : o );
						}
			}
			keys[lane] = v;
		}
		else if( n <= SP_SORTCAP )
		{
			int32_t npow = 64;
			while( npow < n ) npow <<= 1;
			bool unsorted = false;
			for( int32_t i = lane; i < npow; i += 32 )
			{
				const unsigned long long v = i < n ? slice[i] : ~0ull;
				keys[i] = v;
				unsorted = unsorted || ( i + 1 < n && v > slice[i + 1] );
			}
			__syncwarp();
			unsigned uns = __ballot_sync( 0xFFFFFFFFu, unsorted );
			if( uns )
			{
				nsorted++;
				for( int pass = 0; uns && pass < SP_OEPASSES; pass++ )
				{
					bool sw = false;
					for( int ph = 0; ph < 2; ph++ )
					{
						for( int32_t a = 2 * lane + ph; a + 1 < n; a += 64 )
						{
							const unsigned long long va = keys[a], vb = keys[a + 1];
							if( va > vb ) { keys[a] = vb; keys[a + 1] = va; sw = true; }
						}
						__syncwarp();
					}
					uns = __ballot_sync( 0xFFFFFFFFu, sw );
				}
				if( uns )
					rc_sort_smem( keys, npow, lane );
			}
		}
		else
		{
			/* rare: serial insertion sort in place, replayed from HBM */
			if( lane == 0 )
				for( int32_t i = 1; i < n; i++ )
				{
					const unsigned long long key = slice[i];
					int32_t j = i - 1;
					while( j >= 0 && slice[j] > key ) { slice[j + 1] = slice[j]; j--; }
					slice[j + 1] = key;
				}
			sorted = slice;
		}

#ifdef RC_SPANS_DEBUG
		__syncwarp();
		const long long dbg1 = clock64();
#endif
		const int32_t bandix = view * w.nbands + ( line - view * w.height ) / SP_LINES;
		const uint16_t *bsurf = w.bandsurf + (size_t)bandix * ( (size_t)w.max_cands + 2 );
		const int32_t nwords = ( w.bandnloc[bandix] + 32 ) >> 5; /* band-local surface numbers */
		uint32_t *bits = s_bits[warp];
		if( nwords > SP_BITWORDS )
			bits = w.scanbits + (size_t)gwarp * (size_t)w.scanbits_words;
		__syncwarp();

		int32_t nsp = -1;
		const int32_t tail_u = (( x0 + vw_w ) << 20 ) + 0xFFFFF;
		if( sorted == keys && __ballot_sync( 0xFFFFFFFFu, rc_line_has_bmodel( w, view, keys, n, bsurf, lane, 32 )))
		{
			if( lane == 0 )
			{
				int32_t ovf = 0;
				nsp = rc_replay_exact( w, view, line - view * w.height, keys, n, x0, tail_u, bsurf, reinterpret_cast<rc_span_t *>( slice ), &ovf );
				novf += ovf;
				nties++;
			}
			nsp = __shfl_sync( 0xFFFFFFFFu, nsp, 0 );
		}
		else if( sorted == keys )
		{
			/* invalid lanes of a partial round read keys[i >= n]: harmless (never used) */
			if( nwords <= 2 ) nsp = rc_replay_warp<2>( keys, n, x0, tail_u, bsurf, reinterpret_cast<rc_span_t *>( slice ), lane );
			else if( nwords <= 4 ) nsp = rc_replay_warp<4>( keys, n, x0, tail_u, bsurf, reinterpret_cast<rc_span_t *>( slice ), lane );
			else if( nwords <= 8 ) nsp = rc_replay_warp<8>( keys, n, x0, tail_u, bsurf, reinterpret_cast<rc_span_t *>( slice ), lane );
			else if( nwords <= 16 ) nsp = rc_replay_warp<16>( keys, n, x0, tail_u, bsurf, reinterpret_cast<rc_span_t *>( slice ), lane );
		}
		if( nsp < 0 )
		{
			/* degenerate line (or a view listing more faces than the register bit set holds): serial replay */
			for( int32_t k = lane; k < nwords; k += 32 ) bits[k] = 0;
			__syncwarp();
			if( lane == 0 )
			{
				int32_t exc[2 * SP_EXCCAP];
				int32_t ovf = 0;
				nsp = rc_generate_spans_bits( sorted, n, x0, tail_u, bits, nwords, exc, SP_EXCCAP,
					reinterpret_cast<rc_span_t *>( slice ), n + 1, &ovf );
				novf += ovf;
				nserial++;
			}
			nsp = __shfl_sync( 0xFFFFFFFFu, nsp, 0 );
			__syncwarp();
			rc_span_t *sp = reinterpret_cast<rc_span_t *>( slice );
			for( int32_t k = lane; k < nsp; k += 32 ) sp[k].surf = bsurf[sp[k].surf];
		}
		if( lane == 0 )
		{
			w.lineinfo[line].nsp = nsp | RC_LINE_DONE;
			nspans_total += nsp;
		}
		__syncwarp();
#ifdef RC_SPANS_DEBUG
		if( lane == 0 )
		{
			const long long dbg2 = clock64();
			atomicAdd( &w.counters[26], (unsigned long long)( dbg1 - dbg0 )); atomicMax( &w.counters[27], (unsigned long long)( dbg1 - dbg0 ));
			atomicAdd( &w.counters[28], (unsigned long long)( dbg2 - dbg1 )); atomicMax( &w.counters[29], (unsigned long long)( dbg2 - dbg1 ));
			atomicAdd( &w.counters[24], (unsigned long long)n ); atomicMax( &w.counters[25], (unsigned long long)n );
			atomicAdd( &w.counters[30], 1ull );
		}
#endif
	}
#ifdef RC_SPANS_DEBUG
	if( lane == 0 ) atomicMax( &w.counters[31], (unsigned long long)( clock64() - dbg_start ));
#endif
	if( lane == 0 )
	{
		if( nspans_total ) atomicAdd( &w.counters[RC_CTR_SPANS], (unsigned long long)nspans_total );
		if( novf ) atomicAdd( &w.counters[RC_CTR_OVERFLOW_EVENTS], (unsigned long long)novf );
		if( nserial ) atomicAdd( &w.counters[RC_CTR_SERIAL_LINES], (unsigned long long)nserial );
		if( nsorted ) atomicAdd( &w.counters[RC_CTR_SORTED_LINES], (unsigned long long)nsorted );
		if( nties ) atomicAdd( &w.counters[RC_CTR_BMODEL_TIES], (unsigned long long)nties );
	}
}

/*
 * rc_replay_warp for lines whose band numbers more surfaces than a lane can keep in registers (up to 64 words = 2047
 * band-local surfaces).  The set lives in shared memory (sbits) with a 64-bit summary (bit w = word w is not empty); a
 * round of 32 events touches at most 32 different words, and only those need the XOR scan: for every touched word
 * (uniform loop over the group leaders) all lanes scan the toggles that fall into it, which gives every lane that
 * word's state after its own event; the lowest surface outside the touched words comes from the summary.  The rest -
 * who is on top before and after every event, span ends, compaction - is rc_replay_warp's.
 */
__device__ __forceinline__ int32_t rc_replay_sparse( const unsigned long long *keys, int32_t n, int32_t head_iu, int32_t tail_u,
	const uint16_t *bsurf, rc_span_t *spans, uint32_t *sbits, int32_t nwords, int lane )
{
	for( int32_t k = lane; k < nwords; k += 32 ) sbits[k] = 0;
	__syncwarp();
	unsigned long long summ = 0ull;
	int32_t top_carry = 0, last_u = head_iu, nspans = 0;
	const uint32_t lt = ( 1u << lane ) - 1u;
	const int32_t INF = 0x7FFFFFFF;

	for( int32_t c0 = 0; c0 < n; c0 += 32 )
	{
		const int32_t i = c0 + lane;
		const unsigned long long key = i < n ? keys[i] : 0ull;
		const int32_t u = (int32_t)((uint32_t)( key >> 32 ) ^ 0x80000000u );
		const uint32_t both = (uint32_t)key;
		const bool valid = i < n && u <= tail_u;
		const int32_t iu = u >> 20;
		const uint32_t trail = both >> 16, lead = both & 0xFFFFu;
		const bool is_lead = lead != 0;
		const int32_t s = (int32_t)( is_lead ? lead : trail );
		const uint32_t bit = 1u << ( s & 31 );
		const int32_t wsel = valid ? ( s >> 5 ) : -1;
		bool exc = valid && (( trail != 0 && lead != 0 ) || wsel >= nwords );

		/* leaders: the first lane of every group of lanes whose event falls into the same word */
		const uint32_t peers = __match_any_sync( 0xFFFFFFFFu, wsel );
		uint32_t leaders = __ballot_sync( 0xFFFFFFFFu, valid && ( __ffs( (int)peers ) - 1 ) == lane );
		unsigned long long touched = 0ull;
		int32_t top_t = INF;  /* lowest surface among the touched words, after this lane's event */
		uint32_t mine = 0;
		while( leaders )
		{
			const int L = __ffs( (int)leaders ) - 1;
			leaders &= leaders - 1;
			const int32_t wd = __shfl_sync( 0xFFFFFFFFu, wsel, L );
			uint32_t x = wsel == wd ? bit : 0u;
#pragma unroll
			for( int d = 1; d < 32; d <<= 1 )
			{
				const uint32_t o = __shfl_up_sync( 0xFFFFFFFFu, x, d );
				if( lane >= d ) x ^= o;
			}
			x ^= sbits[wd < nwords ? wd : 0]; /* the word after this lane's event (wd >= nwords: the line goes to the serial replay below) */
			if( wsel == wd ) mine = x;
			if( x )
			{
				const int32_t low = ( wd << 5 ) + __ffs( (int)x ) - 1;
				if( low < top_t ) top_t = low;
			}
			touched |= 1ull << wd;
			const uint32_t last = __shfl_sync( 0xFFFFFFFFu, x, 31 );
			__syncwarp();
			if( lane == 0 && wd < nwords ) sbits[wd] = last;
			if( last ) summ |= 1ull << wd; else summ &= ~( 1ull << wd );
		}
		__syncwarp();
		/* lowest surface in the words this round did not touch (their state is the one before the round) */
		int32_t top_u = INF;
		{
			/* summ was updated for the touched words only: the untouched bits are still the old ones */
			const unsigned long long mu = summ & ~touched;
			if( mu )
			{
				const int32_t wu = __ffsll( (long long)mu ) - 1;
				top_u = ( wu << 5 ) + __ffs( (int)sbits[wu] ) - 1;
			}
		}
		int32_t top_after = top_t < top_u ? top_t : top_u;
		if( top_after == INF ) top_after = 0;
		/* lanes past the last valid event: the state after the last valid one */
		/* leading edge: the bit must have been clear (now set); trailing edge: set (now clear) */
		exc = exc || ( valid && ((( mine & bit ) != 0 ) != is_lead ));
		if( __ballot_sync( 0xFFFFFFFFu, exc ))
			return -1;

		int32_t top_before = __shfl_up_sync( 0xFFFFFFFFu, top_after, 1 );
		if( lane == 0 ) top_before = top_carry;
		const bool brk = valid && ( is_lead ? ( top_before == 0 || s < top_before ) : ( s == top_before ));

		int32_t m = brk ? iu : (int32_t)0x80000000;
#pragma unroll
		for( int d = 1; d < 32; d <<= 1 )
		{
			const int32_t o = __shfl_up_sync( 0xFFFFFFFFu, m, d );
			if( lane >= d && o > m ) m = o;
		}
		int32_t prev = __shfl_up_sync( 0xFFFFFFFFu, m, 1 );
		if( lane == 0 || prev < last_u ) prev = last_u;
		const bool emit = brk && iu > prev;
		const uint32_t em = __ballot_sync( 0xFFFFFFFFu, emit );
		if( emit )
		{
			rc_span_t sp;
			sp.surf = bsurf[top_before]; sp.u = (int16_t)prev; sp.count = (int16_t)( iu - prev );
			spans[nspans + __popc( em & lt )] = sp;
		}
		nspans += __popc( em );
		top_carry = __shfl_sync( 0xFFFFFFFFu, top_after, 31 );
		const int32_t mlast = __shfl_sync( 0xFFFFFFFFu, m, 31 );
		if( mlast > last_u ) last_u = mlast;
	}
	const int32_t iu = tail_u >> 20;
	if( iu > last_u )
	{
		if( lane == 0 )
		{
			rc_span_t sp;
			sp.surf = bsurf[top_carry]; sp.u = (int16_t)last_u; sp.count = (int16_t)( iu - last_u );
			spans[nspans] = sp;
		}
		nspans++;
	}
	return nspans;
}

/*
 * The lines k_lines leaves alone (RC_LINE_WIDE: up to WD_SORTCAP events, up to 2047 band-local surfaces - the horizon of
 * a 1080p view down a long row of rooms): one block per line, lines drawn from a cursor over k_events' list.  Keys are
 * sorted in shared memory by the whole block (bitonic network; k_events' presort leaves little to do but an unstaged
 * band arrives unsorted), one warp replays them (rc_replay_sparse); what neither takes is replayed serially.
 */
__global__ void __launch_bounds__( WD_THREADS ) k_lines_wide( rc_work_t w )
{
	__shared__ __align__( 16 ) unsigned long long s_keys[WD_SORTCAP];
	__shared__ uint32_t s_bits[SP_BITWORDS + 1];
	__shared__ int32_t s_item, s_unsorted, s_nsp;
	const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
	const int32_t nwide = *w.wide_count;
	for( ;; )
	{
		__syncthreads();
		if( t == 0 ) { s_item = atomicAdd( w.line_cursor + 2, 1 ); s_unsorted = 0; s_nsp = -1; }
		__syncthreads();
		const int32_t item = s_item;
		if( item >= nwide )
			break;
		const int32_t line = w.wide_list[item];
		const int32_t view = line / w.height;
		const rc_lineinfo_t li = w.lineinfo[line];
		const int32_t n = li.nsp & ~( RC_LINE_WIDE | RC_LINE_DONE );
		if( li.ofs < 0 || ( li.nsp & RC_LINE_DONE ))
			continue;
		const rc_view_t *vw = &w.views[view];
		const int32_t x0 = vw->vrect[0], vw_w = vw->vrect[2];
		unsigned long long *slice = w.events + (size_t)view * w.view_events + li.ofs;
		const int32_t bandix = view * w.nbands + ( line - view * w.height ) / SP_LINES;
		const uint16_t *bsurf = w.bandsurf + (size_t)bandix * ( (size_t)w.max_cands + 2 );
		const int32_t nwords = ( w.bandnloc[bandix] + 32 ) >> 5;
		const int32_t tail_u = (( x0 + vw_w ) << 20 ) + 0xFFFFF;
		const unsigned long long *sorted = s_keys;
		if( n <= WD_SORTCAP )
		{
			int32_t npow = 64;
			while( npow < n ) npow <<= 1;
			bool uns = false;
			for( int32_t i = t; i < npow; i += WD_THREADS )
			{
				const unsigned long long v = i < n ? slice[i] : ~0ull;
				s_keys[i] = v;
				uns = uns || ( i + 1 < n && v > slice[i + 1] );
			}
			if( uns ) s_unsorted = 1;
			__syncthreads();
			if( s_unsorted )
				for( int32_t k = 2; k <= npow; k <<= 1 )
					for( int32_t j = k >> 1; j > 0; j >>= 1 )
					{
						for( int32_t pr = t; pr < ( npow >> 1 ); pr += WD_THREADS )
						{
							const int32_t a = (( pr & ~( j - 1 )) << 1 ) | ( pr & ( j - 1 )), b = a + j;
							const unsigned long long va = s_keys[a], vb = s_keys[b];
							if(( va > vb ) == (( a & k ) == 0 )) { s_keys[a] = vb; s_keys[b] = va; }
						}
						__syncthreads();
					}
		}
		else
		{
			/* beyond the shared-memory sorter: serial insertion sort in place, replayed from HBM */
			if( t == 0 )
				for( int32_t i = 1; i < n; i++ )
				{
					const unsigned long long key = slice[i];
					int32_t j = i - 1;
					while( j >= 0 && slice[j] > key ) { slice[j + 1] = slice[j]; j--; }
					slice[j + 1] = key;
				}
			sorted = slice;
			__syncthreads();
		}
		if( warp == 0 )
		{
			int32_t nsp = -1;
			if( sorted == s_keys && __ballot_sync( 0xFFFFFFFFu, rc_line_has_bmodel( w, view, s_keys, n, bsurf, lane, 32 )))
			{
				if( lane == 0 )
				{
					int32_t ovf = 0;
					nsp = rc_replay_exact( w, view, line - view * w.height, s_keys, n, x0, tail_u, bsurf, reinterpret_cast<rc_span_t *>( slice ), &ovf );
					if( ovf ) atomicAdd( &w.counters[RC_CTR_OVERFLOW_EVENTS], 1ull );
					atomicAdd( &w.counters[RC_CTR_BMODEL_TIES], 1ull );
				}
				nsp = __shfl_sync( 0xFFFFFFFFu, nsp, 0 );
			}
			else if( sorted == s_keys && nwords <= SP_BITWORDS )
				nsp = rc_replay_sparse( s_keys, n, x0, tail_u, bsurf, reinterpret_cast<rc_span_t *>( slice ), s_bits, nwords, lane );
			if( nsp < 0 )
			{
				uint32_t *bits = nwords <= SP_BITWORDS ? s_bits : w.scanbits + ( (size_t)LN_GRID * LN_WARPS - 1 - ( blockIdx.x % ( LN_GRID * LN_WARPS ))) * (size_t)w.scanbits_words;
				for( int32_t k = lane; k < nwords; k += 32 ) bits[k] = 0;
				__syncwarp();
				if( lane == 0 )
				{
					int32_t exc[2 * SP_EXCCAP];
					int32_t ovf = 0;
					nsp = rc_generate_spans_bits( sorted, n, x0, tail_u, bits, nwords, exc, SP_EXCCAP, reinterpret_cast<rc_span_t *>( slice ), n + 1, &ovf );
					if( ovf ) atomicAdd( &w.counters[RC_CTR_OVERFLOW_EVENTS], 1ull );
					atomicAdd( &w.counters[RC_CTR_SERIAL_LINES], 1ull );
				}
				nsp = __shfl_sync( 0xFFFFFFFFu, nsp, 0 );
				__syncwarp();
				rc_span_t *sp = reinterpret_cast<rc_span_t *>( slice );
				for( int32_t k = lane; k < nsp; k += 32 ) sp[k].surf = bsurf[sp[k].surf];
			}
			if( lane == 0 )
			{
				w.lineinfo[line].nsp = nsp | RC_LINE_DONE;
				atomicAdd( &w.counters[RC_CTR_SPANS], (unsigned long long)nsp );
				atomicAdd( &w.counters[RC_CTR_WIDE_LINES], 1ull );
			}
		}
	}
}

#endif /* RC_K_SPANS_CUH */
