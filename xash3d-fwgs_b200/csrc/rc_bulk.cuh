/*
 * rc_bulk.cuh - sm_100a bulk asynchronous copies (the TMA unit's 1-D mode: PTX cp.async.bulk, SASS UBLKCP) and the
 * mbarrier a block waits on for them.  One thread issues a copy of any multiple of 16 bytes between global and shared
 * memory; the copy engine moves it while the block's threads do something else, and reports the bytes on the mbarrier.
 * Used where a block stages a read-only table whole (k_viswalk: the BSP tree) - no reference counterpart.
 */
#ifndef RC_BULK_CUH
#define RC_BULK_CUH

#include <stdint.h>

__device__ __forceinline__ uint32_t rc_smem_addr( const void *p )
{
	return (uint32_t)__cvta_generic_to_shared( p );
}

/* one thread, before anyone uses the barrier; `count` = arrivals per phase */
__device__ __forceinline__ void rc_mbar_init( unsigned long long *bar, uint32_t count )
{
	asm volatile( "mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"( rc_smem_addr( bar )), "r"( count ) : "memory" );
	asm volatile( "fence.mbarrier_init.release.cluster;" ::: "memory" ); /* visible to the async proxy */
}

/* the issuing thread's arrival + the bytes the phase waits for (< 2^20 per phase) */
__device__ __forceinline__ void rc_mbar_expect_tx( unsigned long long *bar, uint32_t bytes )
{
	asm volatile( "mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"( rc_smem_addr( bar )), "r"( bytes ) : "memory" );
}

/* every thread that reads the staged bytes: spin on the phase with the given parity */
__device__ __forceinline__ void rc_mbar_wait( unsigned long long *bar, uint32_t parity )
{
	asm volatile(
		"{\n"
		".reg .pred p;\n"
		"RC_MBAR_WAIT_%=:\n"
		"mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
		"@p bra RC_MBAR_DONE_%=;\n"
		"bra RC_MBAR_WAIT_%=;\n"
		"RC_MBAR_DONE_%=:\n"
		"}\n" :: "r"( rc_smem_addr( bar )), "r"( parity ) : "memory" );
}

/* global -> shared, bytes a multiple of 16, both addresses 16-byte aligned; completion is counted on `bar` */
__device__ __forceinline__ void rc_bulk_g2s( void *dst, const void *src, uint32_t bytes, unsigned long long *bar )
{
	asm volatile( "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
		:: "r"( rc_smem_addr( dst )), "l"( src ), "r"( bytes ), "r"( rc_smem_addr( bar )) : "memory" );
}

/* a table in pieces of RC_BULK_PIECE bytes, starting at piece `rot`: blocks that stage the same table at the same time
 * start at different places, so their requests spread over the L2 slices */
#define RC_BULK_PIECE 8192u
__device__ __forceinline__ void rc_bulk_table_g2s( void *dst, const void *src, uint32_t bytes, unsigned long long *bar, uint32_t rot )
{
	const uint32_t n = ( bytes + RC_BULK_PIECE - 1 ) / RC_BULK_PIECE;
	for( uint32_t i = 0; i < n; i++ )
	{
		uint32_t c = i + rot % n;
		if( c >= n ) c -= n;
		const uint32_t off = c * RC_BULK_PIECE, len = min( RC_BULK_PIECE, bytes - off );
		rc_bulk_g2s( (unsigned char *)dst + off, (const unsigned char *)src + off, len, bar );
	}
}

#endif /* RC_BULK_CUH */
