/*
 * rc_k_spans.cuh - S3a, span generation: R_ScanEdges / R_InsertNewEdges / R_StepActiveU / R_GenerateSpans /
 * R_LeadingEdge / R_TrailingEdge / R_CleanupSpan (ref/soft/r_edge.c:122-675) for every scanline of every view.
 *
 * The reference sweeps a view top to bottom keeping one sorted active-edge list.  An edge's u at scanline y is
 * closed form (u0 + (y - v) * u_step in int32 wraparound, r_edge.c:187), so scanlines are independent given
 * the edge set.  Two kernels:
 *
 *   k_events  block per (view, band of SP_LINES = 32 scanlines), lane = scanline, warps stride over the band's
 *             edge list (binned by k_faces): pass 1 counts the edges crossing every line in registers, a block
 *             prefix sum gives every line its slice of the band's arena in HBM, pass 2 writes the event keys
 *             {u at the line, surfaces}.  No atomics.
 *   k_lines   warp per scanline, lines strided over the whole grid so that the heavy lines of one view land on
 *             different SMs: load the line's keys (coalesced), bitonic sort (registers + shuffles up to 32 events,
 *             shared memory up to SP_SORTCAP) = the order R_InsertNewEdges / R_StepActiveU keep incrementally,
 *             then replay them against the surface bit set (rc_generate_spans_bits) out of shared memory; the spans
 *             overwrite the slice, the count goes to the line table k_draw reads.
 */
#ifndef RC_K_SPANS_CUH
#define RC_K_SPANS_CUH

/* SP_LINES (scanlines per band) is defined by rc_device.cu: k_faces bins edges by band */
#define EV_THREADS  256
#define EV_WARPS    ( EV_THREADS / 32 )
#define LN_WARPS    4
#define SP_SORTCAP  512  /* events per line the shared-memory sorter takes; longer lines fall back to a serial sort */
#define SP_BITWORDS 64   /* shared-memory bit-set words per line (views listing <= 2015 faces), else global scratch */
#define SP_EXCCAP   8

#if SP_LINES != 32
#error "k_events maps lanes to scanlines: SP_LINES must be 32"
#endif

__global__ void __launch_bounds__( EV_THREADS ) k_events( rc_work_t w )
{
	__shared__ int32_t s_wcnt[EV_WARPS][SP_LINES];

	const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
	const int band = blockIdx.x, view = blockIdx.y, nbands = gridDim.x;
	const rc_view_t *vw = &w.views[view];
	const int32_t y0 = vw->vrect[1], vw_h = vw->vrect[3];
	const int32_t y = band * SP_LINES + lane;
	const bool valid = y >= y0 && y < y0 + vw_h && y < w.height;

	int32_t ne = w.bandcnt[view * nbands + band];
	if( ne > w.band_cap ) ne = w.band_cap;
	const rc_edge_t *gedges = w.bandedges + ((size_t)view * nbands + band ) * w.band_cap;
	unsigned long long *arena = w.events + ((size_t)view * nbands + band ) * w.blk_events;

	/* 1. events per line: this warp's share of the edges */
	int32_t cnt = 0;
	for( int32_t k = warp; k < ne; k += EV_WARPS )
	{
		const uint32_t vv = gedges[k].vv;
		cnt += ( valid && y >= (int32_t)( vv & 0xFFFF ) && y <= (int32_t)( vv >> 16 )) ? 1 : 0;
	}
	s_wcnt[warp][lane] = cnt;
	__syncthreads();
	int32_t mybase = 0, n = 0;
	for( int k = 0; k < EV_WARPS; k++ )
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
		rc_lineinfo_t li;
		li.ofs = -1; li.nsp = 0;
		if( valid && fits )
		{
			li.ofs = band * w.blk_events + off; /* index inside the view's arena */
			li.nsp = n;                         /* events for now; k_lines replaces it with the span count */
		}
		else if( valid )
			atomicAdd( &w.counters[RC_CTR_OVERFLOW_EVENTS], 1ull );
		if( y < w.height )
			w.lineinfo[(size_t)view * w.height + y] = li;
	}
	if( !valid || !fits )
		return;

	/* 2. the events */
	int32_t pos = off + mybase;
	for( int32_t k = warp; k < ne; k += EV_WARPS )
	{
		const rc_edge_t e = gedges[k];
		const int32_t v = (int32_t)( e.vv & 0xFFFF );
		if( y >= v && y <= (int32_t)( e.vv >> 16 ))
			arena[pos++] = rc_event_key( (int32_t)((uint32_t)e.u + (uint32_t)( y - v ) * (uint32_t)e.u_step ), e.surfs );
	}
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

__global__ void __launch_bounds__( LN_WARPS * 32 ) k_lines( rc_work_t w )
{
	__shared__ __align__( 16 ) unsigned long long s_keys[LN_WARPS][SP_SORTCAP];
	__shared__ uint32_t s_bits[LN_WARPS][SP_BITWORDS + 1];

	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const int32_t gwarp = blockIdx.x * LN_WARPS + warp, nwarps = gridDim.x * LN_WARPS;
	const int32_t nlines = w.nviews * w.height;
	unsigned long long *keys = s_keys[warp];
	int32_t nspans_total = 0, novf = 0;

	for( int32_t line = gwarp; line < nlines; line += nwarps )
	{
		const int32_t view = line / w.height;
		const rc_lineinfo_t li = w.lineinfo[line];
		if( li.ofs < 0 )
			continue;
		const rc_view_t *vw = &w.views[view];
		const int32_t x0 = vw->vrect[0], vw_w = vw->vrect[2];
		const int32_t n = li.nsp;
		unsigned long long *slice = w.events + (size_t)view * w.view_events + li.ofs;
		const unsigned long long *sorted = keys;

		if( n <= 32 )
		{
			unsigned long long v = lane < n ? slice[lane] : ~0ull;
			for( int k = 2; k <= 32; k <<= 1 )
				for( int j = k >> 1; j > 0; j >>= 1 )
				{
					const unsigned long long o = __shfl_xor_sync( 0xFFFFFFFFu, v, j );
					const bool keepmin = (( lane & k ) == 0 ) == (( lane & j ) == 0 );
					v = keepmin ? ( v < o ? v : o ) : ( v > o ? v : o );
				}
			keys[lane] = v;
		}
		else if( n <= SP_SORTCAP )
		{
			int32_t npow = 64;
			while( npow < n ) npow <<= 1;
			for( int32_t i = lane; i < npow; i += 32 ) keys[i] = i < n ? slice[i] : ~0ull;
			__syncwarp();
			rc_sort_smem( keys, npow, lane );
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

		const int32_t nwords = ( w.ncands[view] + 32 ) >> 5;
		uint32_t *bits = s_bits[warp];
		if( nwords > SP_BITWORDS )
			bits = w.scanbits + (size_t)gwarp * (size_t)w.scanbits_words;
		for( int32_t k = lane; k < nwords; k += 32 ) bits[k] = 0;
		__syncwarp();

		if( lane == 0 )
		{
			int32_t exc[2 * SP_EXCCAP];
			int32_t ovf = 0;
			const int32_t nsp = rc_generate_spans_bits( sorted, n, x0, (( x0 + vw_w ) << 20 ) + 0xFFFFF, bits, nwords, exc, SP_EXCCAP,
				reinterpret_cast<rc_span_t *>( slice ), n + 1, &ovf );
			w.lineinfo[line].nsp = nsp;
			nspans_total += nsp;
			novf += ovf;
		}
		__syncwarp();
	}
	if( lane == 0 )
	{
		if( nspans_total ) atomicAdd( &w.counters[RC_CTR_SPANS], (unsigned long long)nspans_total );
		if( novf ) atomicAdd( &w.counters[RC_CTR_OVERFLOW_EVENTS], (unsigned long long)novf );
	}
}

#endif /* RC_K_SPANS_CUH */
