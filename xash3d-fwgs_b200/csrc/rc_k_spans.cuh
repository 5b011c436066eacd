/*
 * rc_k_spans.cuh - S3a, span generation: R_ScanEdges / R_InsertNewEdges / R_StepActiveU / R_GenerateSpans /
 * R_LeadingEdge / R_TrailingEdge / R_CleanupSpan (ref/soft/r_edge.c:122-675) for every scanline of every view.
 *
 * The reference sweeps a view top to bottom keeping one sorted active-edge list.  An edge's u at scanline y is
 * closed form (u0 + (y - v) * u_step in int32 wraparound, r_edge.c:187), so scanlines are independent given
 * the edge set.  One block takes one band of SP_LINES scanlines of one view and runs four phases:
 *
 *   1. count   thread per EDGE of the band's list (binned by k_faces): one shared-memory atomic per scanline the
 *              edge crosses -> events per line; block prefix sum -> every line's slice of the band's arena in HBM
 *   2. fill    same walk again: event keys {u at the line, surfaces} into the slices
 *   3. sort    warp per LINE: bitonic sort of the line's keys (registers + shuffles up to 32 events, shared
 *              memory up to SP_SORTCAP), the order R_InsertNewEdges / R_StepActiveU keep incrementally
 *   4. replay  thread per LINE: rc_generate_spans_bits against the surface bit set (shared memory); the spans
 *              overwrite the slice, {slice, count} goes to the line table k_draw reads
 */
#ifndef RC_K_SPANS_CUH
#define RC_K_SPANS_CUH

/* SP_LINES (scanlines per band == threads per block) is defined by rc_device.cu */
#define SP_THREADS  128
#define SP_WARPS    ( SP_THREADS / 32 )
#define SP_SORTCAP  512  /* events per line the shared-memory sorter takes; longer lines fall back to a serial sort */
#define SP_BITWORDS 64   /* shared-memory bit-set words per line (views listing <= 2015 faces), else global scratch */
#define SP_EXCCAP   8


/* bitonic sort of one line's keys by one warp; buf: SP_SORTCAP keys of shared memory owned by the warp */
__device__ __forceinline__ void rc_sort_line( unsigned long long *keys, int32_t n, unsigned long long *buf, int lane )
{
	if( n <= 1 )
		return;
	if( n <= 32 )
	{
		unsigned long long v = lane < n ? keys[lane] : ~0ull;
		for( int k = 2; k <= 32; k <<= 1 )
			for( int j = k >> 1; j > 0; j >>= 1 )
			{
				const unsigned long long o = __shfl_xor_sync( 0xFFFFFFFFu, v, j );
				const bool keepmin = (( lane & k ) == 0 ) == (( lane & j ) == 0 );
				v = keepmin ? ( v < o ? v : o ) : ( v > o ? v : o );
			}
		if( lane < n ) keys[lane] = v;
		return;
	}
	if( n > SP_SORTCAP )
	{
		/* rare: serial insertion sort in place */
		if( lane == 0 )
			for( int32_t i = 1; i < n; i++ )
			{
				const unsigned long long key = keys[i];
				int32_t j = i - 1;
				while( j >= 0 && keys[j] > key ) { keys[j + 1] = keys[j]; j--; }
				keys[j + 1] = key;
			}
		return;
	}
	int32_t npow = 64;
	while( npow < n ) npow <<= 1;
	for( int32_t i = lane; i < npow; i += 32 ) buf[i] = i < n ? keys[i] : ~0ull;
	__syncwarp();
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
	for( int32_t i = lane; i < n; i += 32 ) keys[i] = buf[i];
	__syncwarp();
}

__global__ void __launch_bounds__( SP_THREADS ) k_spans( rc_work_t w )
{
	__shared__ __align__( 16 ) unsigned long long s_sort[SP_WARPS * SP_SORTCAP + SP_LINES];
	uint32_t *s_bits = reinterpret_cast<uint32_t *>( s_sort ); /* SP_LINES * ( SP_BITWORDS + 1 ) words: the sort buffers are dead by the replay phase */
	__shared__ int32_t s_cnt[SP_LINES], s_off[SP_LINES + 1], s_cur[SP_LINES];

	const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
	const int band = blockIdx.x, view = blockIdx.y, nbands = gridDim.x;
	const rc_view_t *vw = &w.views[view];
	const int32_t x0 = vw->vrect[0], y0 = vw->vrect[1], vw_w = vw->vrect[2], vw_h = vw->vrect[3];
	const int32_t ylo = max( band * SP_LINES, y0 ), yhi = min( min( band * SP_LINES + SP_LINES, y0 + vw_h ), w.height ) - 1;

	int32_t ne = w.bandcnt[view * nbands + band];
	if( ne > w.band_cap ) ne = w.band_cap;
	const rc_edge_t *gedges = w.bandedges + ((size_t)view * nbands + band ) * w.band_cap;
	unsigned long long *arena = w.events + ((size_t)view * nbands + band ) * w.blk_events;

	if( t < SP_LINES ) { s_cnt[t] = 0; s_cur[t] = 0; }
	__syncthreads();
#ifdef RC_SPANS_DEBUG
	long long dbg_t0 = clock64(), dbg_t1, dbg_t2, dbg_t3, dbg_t4;
#define DBG_MARK( v ) v = clock64();
#else
#define DBG_MARK( v )
#endif

	/* 1. events per line */
	for( int32_t k = t; k < ne; k += SP_THREADS )
	{
		const uint32_t vv = gedges[k].vv;
		const int32_t a = max( (int32_t)( vv & 0xFFFF ), ylo ), b = min( (int32_t)( vv >> 16 ), yhi );
		for( int32_t y = a; y <= b; y++ )
			atomicAdd( &s_cnt[y - band * SP_LINES], 1 );
	}
	__syncthreads();
	if( t < 32 )
	{
		/* every line gets its event count + 1 slots (the spans that replace the events can be one more) */
		const int32_t c0 = s_cnt[t] + 1, c1 = s_cnt[t + 32] + 1;
		const int32_t i0 = rc_warp_incl_scan( c0, lane );
		const int32_t tot0 = __shfl_sync( 0xFFFFFFFFu, i0, 31 );
		const int32_t i1 = rc_warp_incl_scan( c1, lane ) + tot0;
		s_off[t] = i0 - c0; s_off[t + 32] = i1 - c1;
		if( t == 31 ) s_off[SP_LINES] = i1;
	}
	__syncthreads();

	DBG_MARK( dbg_t1 )
	/* 2. the events */
	for( int32_t k = t; k < ne; k += SP_THREADS )
	{
		const rc_edge_t e = gedges[k];
		const int32_t a = max( (int32_t)( e.vv & 0xFFFF ), ylo ), b = min( (int32_t)( e.vv >> 16 ), yhi );
		uint32_t u = (uint32_t)rc_edge_u_at( &e, a );
		for( int32_t y = a; y <= b; y++ )
		{
			const int32_t l = y - band * SP_LINES;
			const int32_t pos = s_off[l] + atomicAdd( &s_cur[l], 1 );
			if( pos < w.blk_events - 1 )
				arena[pos] = rc_event_key( (int32_t)u, e.surfs );
			u += (uint32_t)e.u_step;
		}
	}
	__syncthreads();

	DBG_MARK( dbg_t2 )
	/* 3. sort every line */
	for( int32_t l = warp; l < SP_LINES; l += SP_WARPS )
	{
		int32_t n = s_cnt[l];
		if( s_off[l] + n > w.blk_events - 1 ) n = max( 0, w.blk_events - 1 - s_off[l] );
		rc_sort_line( arena + s_off[l], n, s_sort + warp * SP_SORTCAP, lane );
	}
	__syncthreads();

	DBG_MARK( dbg_t3 )
	/* 4. replay */
	if( t < SP_LINES )
	{
		const int y = band * SP_LINES + t;
		rc_lineinfo_t li;
		li.ofs = 0; li.nsp = 0;
		if( y >= ylo && y <= yhi && s_off[t] < w.blk_events - 1 )
		{
			const int32_t n = s_cnt[t];
			const int32_t neff = min( n, w.blk_events - 1 - s_off[t] );
			unsigned long long *slice = arena + s_off[t];
			const int32_t nwords = ( w.ncands[view] + 32 ) >> 5;
			uint32_t *bits = s_bits + t * ( SP_BITWORDS + 1 );
			if( nwords > SP_BITWORDS )
				bits = w.scanbits + (((size_t)view * nbands + band ) * SP_LINES + t ) * (size_t)w.scanbits_words;
			for( int32_t k = 0; k < nwords; k++ ) bits[k] = 0;
			int32_t exc[2 * SP_EXCCAP];
			int32_t ovf = neff < n ? 1 : 0;
			li.nsp = rc_generate_spans_bits( slice, neff, x0, (( x0 + vw_w ) << 20 ) + 0xFFFFF, bits, nwords, exc, SP_EXCCAP,
				reinterpret_cast<rc_span_t *>( slice ), neff + 1, &ovf );
			li.ofs = band * w.blk_events + s_off[t]; /* index inside the view's arena */
			if( ovf ) atomicAdd( &w.counters[RC_CTR_OVERFLOW_EVENTS], 1ull );
		}
		if( y < w.height )
			w.lineinfo[(size_t)view * w.height + y] = li;
		int32_t tot = li.nsp;
		for( int s = 16; s > 0; s >>= 1 ) tot += __shfl_xor_sync( 0xFFFFFFFFu, tot, s );
		if( lane == 0 && tot ) atomicAdd( &w.counters[RC_CTR_SPANS], (unsigned long long)tot );
	}
#ifdef RC_SPANS_DEBUG
	__syncthreads();
	DBG_MARK( dbg_t4 )
	if( t < SP_LINES )
	{
		atomicAdd( &w.counters[24], (unsigned long long)s_cnt[t] ); atomicMax( &w.counters[25], (unsigned long long)s_cnt[t] );
		if( s_cnt[t] > 32 ) atomicAdd( &w.counters[26], 1ull );
		if( s_cnt[t] > 64 ) atomicAdd( &w.counters[27], 1ull );
		if( s_cnt[t] > 128 ) atomicAdd( &w.counters[28], 1ull );
	}
	if( t == 0 )
	{
		atomicAdd( &w.counters[29], (unsigned long long)ne ); atomicMax( &w.counters[30], (unsigned long long)ne );
		atomicAdd( &w.counters[16], (unsigned long long)( dbg_t1 - dbg_t0 )); atomicMax( &w.counters[17], (unsigned long long)( dbg_t1 - dbg_t0 ));
		atomicAdd( &w.counters[18], (unsigned long long)( dbg_t2 - dbg_t1 )); atomicMax( &w.counters[19], (unsigned long long)( dbg_t2 - dbg_t1 ));
		atomicAdd( &w.counters[20], (unsigned long long)( dbg_t3 - dbg_t2 )); atomicMax( &w.counters[21], (unsigned long long)( dbg_t3 - dbg_t2 ));
		atomicAdd( &w.counters[22], (unsigned long long)( dbg_t4 - dbg_t3 )); atomicMax( &w.counters[23], (unsigned long long)( dbg_t4 - dbg_t3 ));
	}
#endif
}

#endif /* RC_K_SPANS_CUH */
