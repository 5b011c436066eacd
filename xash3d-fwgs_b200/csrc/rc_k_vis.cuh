/*
 * rc_k_vis.cuh - S0, visibility + BSP walk of one view per block:
 *   Mod_PointInLeaf / Mod_FatPVS (engine/common/mod_bmodel.c:1116,1203), R_MarkLeaves (ref/soft/r_main.c:1021),
 *   R_RenderWorld / R_RecursiveWorldNode (ref/soft/r_bsp.c:466-641).
 *
 * The reference recurses front to back and posts faces as it goes; the only things that reach the later stages
 * are WHICH faces are posted, with which clipflags, and in WHAT ORDER (the order is the surface key).  All three
 * are functions of the visited part of the tree, so the walk is done level by level instead of depth first:
 *
 *   P1  top-down, one barrier pair per tree level: every frontier entry is tested (node visframe bit, 4-plane
 *       bbox cull -> reduced clipflags, leaf contents/visframe); a surviving node appends its two children to the
 *       next level, nearer child first.
 *   P2  bottom-up: size of every visited subtree (visited node or leaf = 1).
 *   P3  top-down: in-order position of every visited entry: [near subtree][node][far subtree] - exactly the order
 *       the recursion reaches them in.
 *   P4  leaf marks: surface s is marked when a visited leaf lists it AND that leaf comes before s's node in the
 *       in-order sequence (the reference tests surf->visframe when it posts the node's faces, i.e. after the near
 *       subtree only).
 *   P5  faces per visited node in in-order sequence, block prefix sum, candidates written at their final index.
 *
 * Bit sets (PVS row, node/leaf visframe, surface marks) live in shared memory: BSP30 caps faces, leafs and nodes
 * at 65535 each, so they never exceed 4 x 8 KB.
 */
#ifndef RC_K_VIS_CUH
#define RC_K_VIS_CUH

#define VW_NTHREADS 512 /* measured on the C2 batch with the tree staged in shared memory: 128 threads 75 us, 256 64 us, 512 62 us */
#define VW_MAXLVL  1024
#define VW_PVSCAP  32   /* leaves inside the fat-PVS radius handled by the parallel OR (r_pvs_radius is 0.1) */

/* per-view int32 scratch: 8 arrays of cap entries (entries, sequence, leaf list, clipflags by position) + nodepos[numnodes] */
static inline size_t rc_vis_scratch_ints( int32_t numnodes )
{
	const size_t cap = 2 * (size_t)numnodes + 2;
	return 8 * cap + (size_t)numnodes + 8;
}

/* + per-node / per-leaf frustum flags of the view (phase A) */
static inline size_t rc_vis_scratch_ints2( int32_t numnodes, int32_t numleafs )
{
	return rc_vis_scratch_ints( numnodes ) + (size_t)numnodes + (size_t)numleafs;
}

/* dynamic shared memory: bit sets, then (when they fit, w.vis_stage) a private copy of the BSP tree records and the
 * per-view scratch - the walk is a chain of dependent loads, shared memory cuts every link from ~300 to ~30 cycles */
#define VW_STAGE_TREE    1
#define VW_STAGE_SCRATCH 2
#define VW_STAGE_LISTS   4 /* surfnode + marksurfaces as u16 */

static inline size_t rc_vis_tree_bytes( const rc_scene_t *sc )
{
	return sizeof( rc_node_t ) * (size_t)sc->numnodes + sizeof( rc_plane_t ) * (size_t)sc->numplanes + sizeof( rc_leaf_t ) * (size_t)sc->numleafs;
}

static_assert( sizeof( rc_node_t ) % 16 == 0 && sizeof( rc_plane_t ) % 16 == 0 && sizeof( rc_leaf_t ) % 16 == 0,
	"the tree's tables are staged with 16-byte bulk copies" );

/* frustum tests of one bbox against all four planes, no early out: bits 0..3 = trivially accepted, 4..7 = rejected */
__device__ __forceinline__ int32_t rc_box_flags( const rc_view_t *vw, const float *minmaxs )
{
	int32_t f = 0;
#pragma unroll
	for( int i = 0; i < 4; i++ )
	{
		const int32_t *pindex = vw->frustum_idx[i];
		const float *nrm = vw->clip_normal[i];
		const float dist = vw->clip_dist[i];
		const float rj = minmaxs[pindex[0]] * nrm[0] + minmaxs[pindex[1]] * nrm[1] + minmaxs[pindex[2]] * nrm[2];
		const float ac = minmaxs[pindex[3]] * nrm[0] + minmaxs[pindex[4]] * nrm[1] + minmaxs[pindex[5]] * nrm[2];
		if( rj <= dist ) f |= 16 << i;
		if( ac >= dist ) f |= 1 << i;
	}
	return f;
}

/* rc_cull_box with float compares: ( (double)dot - (double)dist ) has the sign of ( dot - dist ), so "d <= 0" is
 * "dot <= dist" and "d >= 0" is "dot >= dist" (r_bsp.c:483-520) */
__device__ __forceinline__ int32_t rc_cull_box_f( const rc_view_t *vw, const float *minmaxs, int32_t clipflags )
{
#pragma unroll
	for( int i = 0; i < 4; i++ )
	{
		if( !( clipflags & ( 1 << i )))
			continue;
		const int32_t *pindex = vw->frustum_idx[i];
		const float *nrm = vw->clip_normal[i];
		const float dist = vw->clip_dist[i];
		const float rj = minmaxs[pindex[0]] * nrm[0] + minmaxs[pindex[1]] * nrm[1] + minmaxs[pindex[2]] * nrm[2];
		if( rj <= dist )
			return -1;
		const float ac = minmaxs[pindex[3]] * nrm[0] + minmaxs[pindex[4]] * nrm[1] + minmaxs[pindex[5]] * nrm[2];
		if( ac >= dist )
			clipflags &= ~( 1 << i );
	}
	return clipflags;
}

template <int VW_THREADS>
__global__ void __launch_bounds__( VW_THREADS ) k_viswalk( rc_scene_t sc, rc_work_t w )
{
	extern __shared__ __align__( 16 ) uint32_t s_bits[];
	__shared__ int32_t s_lvl[VW_MAXLVL + 2];
	__shared__ int32_t s_n, s_nleaf, s_npvs, s_allvis, s_ovf;
	__shared__ int32_t s_pvsleaf[VW_PVSCAP];
	__shared__ int32_t s_warp[VW_THREADS / 32];
	__shared__ rc_view_t s_view;
	__shared__ __align__( 8 ) unsigned long long s_bar; /* mbarrier of the tree's bulk copies */

	const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
#ifdef RC_SPANS_DEBUG
	long long dbg_t[10]; int dbg_n = 0;
#define VW_MARK() do { __syncthreads(); dbg_t[dbg_n++] = clock64(); } while( 0 )
#else
#define VW_MARK() do { } while( 0 )
#endif
	const rc_view_t *vw = &s_view;
	const int32_t viswords = w.visbytes_padded >> 2;
	uint32_t *vis = s_bits;
	uint32_t *nodevis = vis + viswords;
	uint32_t *leafmark = nodevis + w.nodewords;
	uint32_t *surfmark = leafmark + w.leafwords;
	const int32_t allwords = viswords + w.nodewords + w.leafwords + w.surfwords;

	const int32_t cap = 2 * sc.numnodes + 2;
	/* the block is persistent over views (grid = resident blocks): the tree is staged and the scratch set up once */
	int32_t *scr = w.walkstack + (size_t)blockIdx.x * w.stack_ints;
	const uint16_t *sn16 = nullptr, *mk16 = nullptr;
	{
		unsigned char *p = reinterpret_cast<unsigned char *>( s_bits ) + (((size_t)allwords * 4 + 15 ) & ~(size_t)15 );
		const bool tree = ( w.vis_stage & VW_STAGE_TREE ) != 0;
		if( tree )
		{
			/* nodes, planes, leafs: bulk asynchronous copies issued by one thread (rc_bulk.cuh); the block goes on
			 * with the 16-bit lists below while the bytes arrive, and waits on the mbarrier before the first view */
			const uint32_t nb = (uint32_t)( sizeof( rc_node_t ) * (size_t)sc.numnodes ), pb = (uint32_t)( sizeof( rc_plane_t ) * (size_t)sc.numplanes ),
				lb = (uint32_t)( sizeof( rc_leaf_t ) * (size_t)sc.numleafs );
			if( t == 0 )
			{
				rc_mbar_init( &s_bar, 1 );
				rc_mbar_expect_tx( &s_bar, nb + pb + lb );
				rc_bulk_table_g2s( p, sc.nodes, nb, &s_bar, blockIdx.x );
				rc_bulk_table_g2s( p + nb, sc.planes, pb, &s_bar, blockIdx.x );
				rc_bulk_table_g2s( p + nb + pb, sc.leafs, lb, &s_bar, blockIdx.x );
			}
			sc.nodes = reinterpret_cast<const rc_node_t *>( p );
			sc.planes = reinterpret_cast<const rc_plane_t *>( p + nb );
			sc.leafs = reinterpret_cast<const rc_leaf_t *>( p + nb + pb );
			p += nb + pb + lb;
		}
		if( w.vis_stage & VW_STAGE_SCRATCH )
		{
			scr = reinterpret_cast<int32_t *>( p ); p += sizeof( int32_t ) * (size_t)w.stack_ints;
		}
		if( w.vis_stage & VW_STAGE_LISTS )
		{
			/* BSP30 caps nodes and faces at 65535: 16 bits are enough */
			uint16_t *a = reinterpret_cast<uint16_t *>( p ), *b = a + sc.numsurfaces;
			for( int i = t; i < sc.numsurfaces; i += VW_THREADS ) a[i] = (uint16_t)sc.surfnode[i];
			for( int i = t; i < sc.nummarks; i += VW_THREADS ) b[i] = (uint16_t)sc.marksurfaces[i];
			sn16 = a; mk16 = b;
		}
		if( tree )
		{
			__syncthreads(); /* the barrier's initialisation is visible to every waiter */
			rc_mbar_wait( &s_bar, 0 );
		}
	}
	int32_t *e_ref = scr, *e_info = scr + cap, *e_child = scr + 2 * (size_t)cap, *e_size = scr + 3 * (size_t)cap;
	int32_t *e_pos = scr + 4 * (size_t)cap, *seq = scr + 5 * (size_t)cap, *leaflist = scr + 6 * (size_t)cap;
	int32_t *posclip = scr + 7 * (size_t)cap;
	int32_t *nodepos = scr + 8 * (size_t)cap;
	int32_t *nflag = nodepos + sc.numnodes + 8, *lflag = nflag + sc.numnodes;

	for( int view = blockIdx.x; view < w.nviews; view += gridDim.x )
	{
#ifdef RC_SPANS_DEBUG
	dbg_n = 0;
#endif
	__syncthreads(); /* the previous view is done with s_view and the scratch */
	VW_MARK();
	if( t < (int)( sizeof( rc_view_t ) / 4 ))
		reinterpret_cast<uint32_t *>( &s_view )[t] = reinterpret_cast<const uint32_t *>( &w.views[view] )[t];
	for( int i = t; i < allwords; i += VW_THREADS ) s_bits[i] = 0;
	for( int i = t; i < sc.numnodes; i += VW_THREADS ) nodepos[i] = -1; /* not visited: precedes no leaf */
	__syncthreads();
	VW_MARK(); /* 1: staged */
	if( t == 0 )
	{
		s_n = 1; s_nleaf = 0; s_npvs = 0; s_allvis = 0; s_ovf = 0;
		e_ref[0] = sc.headnode; e_info[0] = 15;

		/* Mod_FatPVS: the leaves within pvs_radius of the eye (front side first, back side deferred) */
		if( vw->pvs_radius < 0 )
			s_allvis = 1;
		else
		{
			/* one walk for Mod_PointInLeaf and Mod_FatPVS: where both sides are within the radius the walk follows
			 * the side the eye is on (d <= 0: back) and defers the other, so the first leaf reached is the view leaf */
			if( !sc.visdata_len )
				s_allvis = 1;
			else
			{
				int32_t *stack = seq; /* free until P5 */
				int32_t sp = 0, node = sc.headnode, np = 0;
				bool first = true;
				for( ;; )
				{
					while( node >= 0 )
					{
						const rc_node_t *nd = &sc.nodes[node];
						const float d = rc_planediff( vw->origin, &sc.planes[nd->plane] );
						if( d > vw->pvs_radius ) node = nd->child[0];
						else if( d < -vw->pvs_radius ) node = nd->child[1];
						else
						{
							const int near = d <= 0 ? 1 : 0;
							if( sp < cap ) stack[sp++] = nd->child[!near];
							node = nd->child[near];
						}
					}
					const int32_t li = -1 - node;
					if( first && sc.leafs[li].cluster < 0 )
					{
						s_allvis = 1;
						break;
					}
					first = false;
					if( sc.leafs[li].cluster >= 0 )
					{
						if( np < VW_PVSCAP ) s_pvsleaf[np++] = li;
						else
						{
							/* rare: more leaves than the list holds - OR this row right here */
							const uint32_t *row = sc.pvsdense + (size_t)li * viswords;
							for( int32_t k = 0; k < viswords; k++ ) vis[k] |= row[k];
						}
					}
					if( !sp ) break;
					node = stack[--sp];
				}
				s_npvs = np;
			}
		}
	}
	__syncthreads();

	/* fat PVS = OR of the (pre-expanded) rows */
	if( s_allvis )
	{
		for( int i = t; i < viswords; i += VW_THREADS ) vis[i] = 0xFFFFFFFFu;
	}
	else
	{
		const int32_t np = s_npvs;
		for( int i = t; i < viswords; i += VW_THREADS )
		{
			uint32_t v = vis[i];
			for( int32_t k = 0; k < np; k++ ) v |= sc.pvsdense[(size_t)s_pvsleaf[k] * viswords + i];
			vis[i] = v;
		}
	}
	__syncthreads();

	VW_MARK(); /* 2: fat pvs */
	/* R_MarkLeaves: threads take clusters; a thread stops climbing where another one already passed */
	for( int i = t; i < sc.numclusters; i += VW_THREADS )
	{
		if( !(( vis[i >> 5] >> ( i & 31 )) & 1 ))
			continue;
		atomicOr( &leafmark[( i + 1 ) >> 5], 1u << (( i + 1 ) & 31 ));
		int32_t node = sc.leafs[i + 1].parent;
		while( node >= 0 )
		{
			const uint32_t bit = 1u << ( node & 31 );
			if( atomicOr( &nodevis[node >> 5], bit ) & bit )
				break;
			node = sc.nodes[node].parent;
		}
	}
	__syncthreads();

	/* phase A: everything about a visible node / leaf that depends on the view only - frustum flags of its bbox,
	 * the side of its plane the eye is on, which of its faces look at the eye - for all of them at once, so that the
	 * level-by-level descent below is bit logic */
	for( int i = t; i < sc.numnodes; i += VW_THREADS )
	{
		if( !rc_testbit( nodevis, i ))
			continue;
		const rc_node_t *node = &sc.nodes[i];
		const double dot = rc_node_dot( vw, &sc.planes[node->plane] );
		const int32_t want = dot < -RC_BACKFACE_EPSILON ? 1 : ( dot > RC_BACKFACE_EPSILON ? 0 : 2 );
		nflag[i] = rc_box_flags( vw, node->minmaxs ) | (( dot >= 0 ? 0 : 1 ) << 8 ) | ( want << 9 );
	}
	for( int i = t; i < sc.numleafs; i += VW_THREADS )
	{
		if( !rc_testbit( leafmark, i ))
			continue;
		const rc_leaf_t *leaf = &sc.leafs[i];
		lflag[i] = leaf->contents == RC_CONTENTS_SOLID ? 0xF0 | 256 : rc_box_flags( vw, leaf->minmaxs );
	}
	__syncthreads();
	VW_MARK(); /* 3: mark leaves */
	/* P1: level-synchronous descent */
	int32_t nlvl = 0;
	if( t == 0 ) s_lvl[0] = 0;
	for( ;; )
	{
		__syncthreads();
		const int32_t a = s_lvl[nlvl], b = s_n;
		__syncthreads();
		if( a == b )
			break;
		if( nlvl + 1 > VW_MAXLVL )
		{
			if( t == 0 ) s_ovf = 1;
			break;
		}
		if( t == 0 ) s_lvl[nlvl + 1] = b;
		for( int32_t e = a + t; e < b; e += VW_THREADS )
		{
			const int32_t ref = e_ref[e];
			int32_t clipflags = e_info[e] & 15, alive = 0, child = -1;
			if( ref < 0 )
			{
				const int32_t li = -1 - ref;
				if( rc_testbit( leafmark, li ))
				{
					const int32_t f = lflag[li];
					if( !( f & 256 ) && !(( f >> 4 ) & clipflags ))
					{
						alive = 1;
						leaflist[atomicAdd( &s_nleaf, 1 )] = e;
					}
				}
			}
			else if( rc_testbit( nodevis, ref ))
			{
				/* R_RecursiveWorldNode tests the planes in order and stops at the first rejection; a plane is only
				 * tested while its flag is still set, and flags only ever get cleared: the outcome is the same as
				 * "any still-flagged plane rejects" */
				const int32_t f = nflag[ref];
				if( !(( f >> 4 ) & clipflags ))
				{
					alive = 1;
					clipflags &= ~f & 15;
					const int side = ( f >> 8 ) & 1;
					const rc_node_t *node = &sc.nodes[ref];
					child = atomicAdd( &s_n, 2 );
					e_ref[child] = node->child[side]; e_info[child] = clipflags;
					e_ref[child + 1] = node->child[!side]; e_info[child + 1] = clipflags;
				}
				else clipflags = 0;
			}
			e_info[e] = clipflags | ( alive << 4 );
			e_child[e] = child;
		}
		nlvl++;
	}
	__syncthreads();

	VW_MARK(); /* 4: P1 */
	/* P2: subtree sizes, deepest level first */
	for( int32_t L = nlvl - 1; L >= 0; L-- )
	{
		const int32_t a = s_lvl[L], b = s_lvl[L + 1];
		for( int32_t e = a + t; e < b; e += VW_THREADS )
		{
			int32_t sz = 0;
			if( e_info[e] & 16 )
			{
				const int32_t c = e_child[e];
				sz = 1 + ( c >= 0 ? e_size[c] + e_size[c + 1] : 0 );
			}
			e_size[e] = sz;
		}
		__syncthreads();
	}

	VW_MARK(); /* 5: P2 */
	/* P3: in-order positions */
	if( t == 0 ) e_pos[0] = 0;
	__syncthreads();
	for( int32_t L = 0; L < nlvl; L++ )
	{
		const int32_t a = s_lvl[L], b = s_lvl[L + 1];
		for( int32_t e = a + t; e < b; e += VW_THREADS )
		{
			if( !( e_info[e] & 16 ))
				continue;
			const int32_t c = e_child[e], base = e_pos[e];
			if( c >= 0 )
			{
				const int32_t mine = base + e_size[c];
				e_pos[c] = base;
				e_pos[c + 1] = mine + 1;
				e_pos[e] = mine;
				nodepos[e_ref[e]] = mine;
				posclip[mine] = e_info[e] & 15;
			}
		}
		__syncthreads();
	}

	VW_MARK(); /* 6: P3 */
	/* P4: leaf marks that precede the surface's node */
	{
		const int32_t nleaf = s_nleaf;
		for( int32_t k = warp; k < nleaf; k += VW_THREADS / 32 )
		{
			const int32_t e = leaflist[k];
			const int32_t pos = e_pos[e];
			const rc_leaf_t *leaf = &sc.leafs[-1 - e_ref[e]];
			const int32_t m0 = leaf->firstmark, m1 = m0 + leaf->nummark;
			for( int32_t m = m0 + lane; m < m1; m += 32 )
			{
				const int32_t s = mk16 ? (int32_t)mk16[m] : sc.marksurfaces[m];
				if( pos < nodepos[sn16 ? (int32_t)sn16[s] : sc.surfnode[s]] )
					atomicOr( &surfmark[s >> 5], 1u << ( s & 31 ));
			}
		}
	}
	__syncthreads();

	VW_MARK(); /* 7: P4 */
	/* P5: the marked, viewer-facing surfaces are the candidates; their order is (in-order position of their
	 * node, surface index).  Threads walk the set bits of the mark set: count per position, block prefix sum,
	 * then every candidate lands at position base + its rank among the node's candidates. */
	const int32_t nseq = nlvl ? e_size[0] : 0;
	for( int32_t i = t; i < nseq; i += VW_THREADS ) seq[i] = 0;
	__syncthreads();
	/* brush entities: a visited leaf's place in the sequence is its key (LEAF_KEY, r_bsp.c:541); it reserves a surface slot
	 * there for every face of every entity whose box touches the leaf - k_bmodels posts the face pieces into them */
	if( w.nbents )
	{
		const int32_t nleaf = s_nleaf;
		for( int32_t k = t; k < nleaf; k += VW_THREADS )
		{
			const int32_t e = leaflist[k];
			const int32_t n = rc_leaf_bmodel_slots( &sc.leafs[-1 - e_ref[e]], w.bents, w.nbents, w.nbents );
			if( n ) seq[e_pos[e]] = n;
		}
	}
	/* thread per surface (a word of the mark set per thread would leave most of the block idle and walk up to 32
	 * bits serially); a thread only ever changes its own bit */
	for( int32_t s = t; s < sc.numsurfaces; s += VW_THREADS )
	{
		const int32_t wi = s >> 5;
		if( !(( surfmark[wi] >> ( s & 31 )) & 1 ))
			continue;
		const int32_t n = sn16 ? (int32_t)sn16[s] : sc.surfnode[s];
		const int want = ( nflag[n] >> 9 ) & 3; /* 2: the eye is on the plane, nothing is posted */
		const int back = ( sc.surfback[wi] >> ( s & 31 )) & 1;
		if( back == want )
			atomicAdd( &seq[nodepos[n]], 1 );
		else
			atomicAnd( &surfmark[wi], ~( 1u << ( s & 31 ))); /* marked but facing away: not a candidate */
	}
	__syncthreads();
	int32_t carry = 0;
	for( int32_t i0 = 0; i0 < nseq; i0 += VW_THREADS )
	{
		const int32_t i = i0 + t;
		const int32_t v = i < nseq ? seq[i] : 0;
		int32_t x = rc_warp_incl_scan( v, lane );
		if( lane == 31 ) s_warp[warp] = x;
		__syncthreads();
		int32_t wbase = 0, tot = 0;
		for( int k = 0; k < VW_THREADS / 32; k++ )
		{
			if( k < warp ) wbase += s_warp[k];
			tot += s_warp[k];
		}
		if( i < nseq ) seq[i] = carry + wbase + x - v;
		carry += tot;
		__syncthreads();
	}
	const int32_t total = carry;
	rc_cand_t *cands = w.cands + (size_t)view * w.max_cands;
	if( w.nbents )
	{
		const int32_t nleaf = s_nleaf;
		int32_t *leafbase = w.leafbase + (size_t)view * sc.numleafs;
		for( int32_t k = t; k < nleaf; k += VW_THREADS )
		{
			const int32_t e = leaflist[k], leaf = -1 - e_ref[e], pos = e_pos[e];
			const int32_t base = seq[pos], n = ( pos + 1 < nseq ? seq[pos + 1] : total ) - base;
			leafbase[leaf] = base;
			if( w.leafkey ) w.leafkey[(size_t)view * sc.numleafs + leaf] = pos; /* the leaf's place in the walk: translucent brush entities key their pieces by it */
			for( int32_t o = base; o < base + n && o < w.max_cands; o++ ) { cands[o].surf = -1; cands[o].clipflags = 0; }
		}
		/* R_FindTopnode and R_RecursiveClipBPoly test node / leaf visframe: the PVS marks leave the block */
		uint32_t *out = w.bmvis + (size_t)view * ( w.nodewords + w.leafwords );
		for( int32_t i = t; i < w.nodewords + w.leafwords; i += VW_THREADS ) out[i] = nodevis[i]; /* leafmark follows nodevis */
	}
	uint16_t *candof = w.candof + (size_t)view * ( sc.numsurfaces + 2 );
	for( int32_t s = t; s < sc.numsurfaces; s += VW_THREADS )
	{
		const int32_t wi = s >> 5, b = s & 31;
		if( !(( surfmark[wi] >> b ) & 1 ))
		{
			candof[s] = 0;
			continue;
		}
		const int32_t n = sn16 ? (int32_t)sn16[s] : sc.surfnode[s], pos = nodepos[n];
		/* rank among the node's candidates: surviving marks in [firstsurface, s) */
		const int32_t f = sc.nodes[n].firstsurface;
		int32_t rank = 0;
		for( int32_t k = f >> 5; k <= wi; k++ )
		{
			uint32_t mk = surfmark[k];
			if( k == ( f >> 5 )) mk &= ~0u << ( f & 31 );
			if( k == wi ) mk &= ( 1u << b ) - 1u;
			rank += __popc( mk );
		}
		const int32_t o = seq[pos] + rank;
		candof[s] = o < w.max_cands ? (uint16_t)( o + 1 ) : (uint16_t)0;
		if( o < w.max_cands )
		{
			cands[o].surf = s;
			cands[o].clipflags = posclip[pos];
		}
	}

	VW_MARK(); /* 8: P5 */
#ifdef RC_SPANS_DEBUG
	if( t == 0 )
		for( int k = 0; k < 8; k++ )
		{
			atomicAdd( &w.counters[32 + 2 * k], (unsigned long long)( dbg_t[k + 1] - dbg_t[k] ));
			atomicMax( &w.counters[33 + 2 * k], (unsigned long long)( dbg_t[k + 1] - dbg_t[k] ));
		}
#endif
	if( t == 0 )
	{
		w.ncands[view] = total < w.max_cands ? total : w.max_cands;
		w.nedges[view] = 0;
		atomicAdd( &w.counters[RC_CTR_CANDS], (unsigned long long)( total < w.max_cands ? total : w.max_cands ));
		atomicAdd( &w.counters[RC_CTR_VIEWS], 1ull );
		atomicAdd( &w.counters[RC_CTR_PIXELS], (unsigned long long)vw->vrect[2] * vw->vrect[3] );
		if( s_ovf ) atomicAdd( &w.counters[RC_CTR_OVERFLOW_STACK], 1ull );
		if( total > w.max_cands ) atomicAdd( &w.counters[RC_CTR_OVERFLOW_CANDS], 1ull );
		atomicMax( &w.counters[RC_CTR_MAX_CANDS], (unsigned long long)total );
	}
	} /* views */
}

#endif /* RC_K_VIS_CUH */
