/*
 * rc_tri.h - the triangle rasteriser of ref_soft restated as scalar __host__ __device__ functions:
 * R_SetupFinalVert / R_AliasProjectAndClipTestFinalVert (ref/soft/r_trialias.c:123-169), R_AliasClipTriangle and the
 * five clippers (r_aclip.c:44-330), R_DrawTriangle / R_PolysetSetEdgeTable / R_PolysetCalcGradients /
 * R_RasterizeAliasPolySmooth / R_PolysetScanLeftEdge_C (r_polyse.c:107-720) and the span drawers
 * R_PolysetFillSpans8 / ...Additive / ...Glow / ...TextureBlended / ...Blended (r_polyse.c:724-1214).
 *
 * The reference walks a triangle scanline by scanline carrying an integer DDA on its left and right edges.  Every
 * scanline's state is closed form in the number of steps taken (the number of "extra" steps after k rows of an edge
 * with error terms (up, down) is floor((k * up - 1) / down) + 1), so rows - and pixels inside a row - can be evaluated
 * independently and in any order with bit-identical results: rc_tri_row gives the left pixel, the pixel count and the
 * s, t, light, 1/z accumulators of a row, rc_tri_pixel the texel index, light and 1/z of a pixel in it.
 */
#ifndef RC_TRI_H
#define RC_TRI_H

#include "rc_core.h"

#define RC_ALIAS_LEFT_CLIP   0x0001 /* ref/soft/r_local.h:608-612 */
#define RC_ALIAS_TOP_CLIP    0x0002
#define RC_ALIAS_RIGHT_CLIP  0x0004
#define RC_ALIAS_BOTTOM_CLIP 0x0008
#define RC_ALIAS_Z_CLIP      0x0010
#define RC_ALIAS_Z_CLIP_PLANE 4.0f

typedef struct
{
	int32_t u, v, s, t, l, zi;
	int32_t flags;
	float   xyz[3];
} rc_fv_t; /* finalvert_t, r_local.h:660 */

/* R_AliasProjectAndClipTestFinalVert, r_trialias.c:123 (r_aliasuvscale is 1: the alias rectangle is the view rectangle) */
RC_HD void rc_fv_project( const rc_view_t *vw, rc_fv_t *fv )
{
	const float x = fv->xyz[0], y = fv->xyz[1], z = fv->xyz[2];
	const float zi = 1.0f / z;
	fv->zi = rc_f2i( zi * ((float)0x8000 * (float)0x10000 ));
	fv->u = rc_f2i(( x * vw->xscale * zi ) + vw->xcenter );
	fv->v = rc_f2i(( y * vw->yscale * zi ) + vw->ycenter );
	if( fv->u < vw->vrect[0] ) fv->flags |= RC_ALIAS_LEFT_CLIP;
	if( fv->v < vw->vrect[1] ) fv->flags |= RC_ALIAS_TOP_CLIP;
	if( fv->u > vw->vrect[0] + vw->vrect[2] ) fv->flags |= RC_ALIAS_RIGHT_CLIP;
	if( fv->v > vw->vrect[1] + vw->vrect[3] ) fv->flags |= RC_ALIAS_BOTTOM_CLIP;
}

/* R_SetupFinalVert, r_trialias.c:146: world point -> view space (aliastransform), near test, projection */
RC_HD void rc_fv_setup( const rc_view_t *vw, rc_fv_t *fv, float x, float y, float z, int32_t light, int32_t s, int32_t t )
{
	const float v[3] = { x, y, z };
	fv->xyz[0] = rc_dot3( v, vw->aliastransform[0] ) + vw->aliastransform[0][3];
	fv->xyz[1] = rc_dot3( v, vw->aliastransform[1] ) + vw->aliastransform[1][3];
	fv->xyz[2] = rc_dot3( v, vw->aliastransform[2] ) + vw->aliastransform[2][3];
	fv->flags = 0;
	fv->l = light;
	fv->u = fv->v = fv->zi = 0;
	if( fv->xyz[2] < RC_ALIAS_Z_CLIP_PLANE )
		fv->flags |= RC_ALIAS_Z_CLIP;
	else
		rc_fv_project( vw, fv );
	fv->s = (int32_t)((uint32_t)s << 16 );
	fv->t = (int32_t)((uint32_t)t << 16 );
}

/* ------------------------------------------------------------------------------------------ */
/* clipping, r_aclip.c                                                                          */

/* int + ( int - int ) * float, truncated on assignment: the expression the clippers are made of */
RC_HD int32_t rc_lerp_i( int32_t a, int32_t b, float scale, float bias )
{
	return rc_f2i((float)a + (float)( b - a ) * scale + bias );
}

/* R_Alias_clip_z, r_aclip.c:44: pfv0 is the unclipped vertex, pfv1 the z-clipped one */
RC_HD void rc_clip_z( const rc_view_t *vw, const rc_fv_t *p0, const rc_fv_t *p1, rc_fv_t *out )
{
	const float scale = ( RC_ALIAS_Z_CLIP_PLANE - p0->xyz[2] ) / ( p1->xyz[2] - p0->xyz[2] );
	out->xyz[0] = p0->xyz[0] + ( p1->xyz[0] - p0->xyz[0] ) * scale;
	out->xyz[1] = p0->xyz[1] + ( p1->xyz[1] - p0->xyz[1] ) * scale;
	out->xyz[2] = RC_ALIAS_Z_CLIP_PLANE;
	out->s = rc_lerp_i( p0->s, p1->s, scale, 0.0f );
	out->t = rc_lerp_i( p0->t, p1->t, scale, 0.0f );
	out->l = rc_lerp_i( p0->l, p1->l, scale, 0.0f );
	/* out->flags is whatever the ping-pong buffer held; the caller overwrites it (r_aclip.c:190) */
	rc_fv_project( vw, out );
}

/* R_Alias_clip_left / right / top / bottom, r_aclip.c:62-173: the edge is walked from its lower vertex (larger v) */
RC_HD void rc_clip_screen( const rc_fv_t *p0, const rc_fv_t *p1, rc_fv_t *out, int horizontal, int32_t bound )
{
	const rc_fv_t *a = p0, *b = p1;
	if( !( p0->v >= p1->v )) { a = p1; b = p0; }
	const float scale = horizontal ? (float)( bound - a->u ) / (float)( b->u - a->u ) : (float)( bound - a->v ) / (float)( b->v - a->v );
	out->u = rc_lerp_i( a->u, b->u, scale, 0.5f );
	out->v = rc_lerp_i( a->v, b->v, scale, 0.5f );
	out->s = rc_lerp_i( a->s, b->s, scale, 0.5f );
	out->t = rc_lerp_i( a->t, b->t, scale, 0.5f );
	out->l = rc_lerp_i( a->l, b->l, scale, 0.5f );
	out->zi = rc_lerp_i( a->zi, b->zi, scale, 0.5f );
}

/* R_AliasClip, r_aclip.c:175; which: 0 z, 1 left, 2 right, 3 bottom, 4 top */
RC_HD int rc_alias_clip( const rc_view_t *vw, const rc_fv_t *in, rc_fv_t *out, int flag, int count, int which )
{
	const int32_t x0 = vw->vrect[0], y0 = vw->vrect[1], x1 = vw->vrect[0] + vw->vrect[2], y1 = vw->vrect[1] + vw->vrect[3];
	int j = count - 1, k = 0;
	for( int i = 0; i < count; j = i, i++ )
	{
		const int oldflags = in[j].flags & flag, flags = in[i].flags & flag;
		if( flags && oldflags )
			continue;
		if( oldflags ^ flags )
		{
			/* fields a clipper does not assign keep what the buffer held: start from the slot's previous content is not
			 * reproducible, but every field the rasteriser reads is assigned below */
			switch( which )
			{
			case 0: rc_clip_z( vw, &in[j], &in[i], &out[k] ); break;
			case 1: rc_clip_screen( &in[j], &in[i], &out[k], 1, x0 ); break;
			case 2: rc_clip_screen( &in[j], &in[i], &out[k], 1, x1 ); break;
			case 3: rc_clip_screen( &in[j], &in[i], &out[k], 0, y1 ); break;
			default: rc_clip_screen( &in[j], &in[i], &out[k], 0, y0 ); break;
			}
			out[k].flags = 0;
			if( out[k].u < x0 ) out[k].flags |= RC_ALIAS_LEFT_CLIP;
			if( out[k].v < y0 ) out[k].flags |= RC_ALIAS_TOP_CLIP;
			if( out[k].u > x1 ) out[k].flags |= RC_ALIAS_RIGHT_CLIP;
			if( out[k].v > y1 ) out[k].flags |= RC_ALIAS_BOTTOM_CLIP;
			k++;
		}
		if( !flags )
		{
			out[k] = in[i];
			k++;
		}
	}
	return k;
}

/* R_AliasClipTriangle, r_aclip.c:217: -> polygon of k vertices in poly (k <= 8), 0 = nothing left.  The triangles
 * the reference then draws are the fan (0, i, i + 1), i = 1 .. k - 2. */
RC_HD int rc_alias_clip_triangle( const rc_view_t *vw, const rc_fv_t *i0, const rc_fv_t *i1, const rc_fv_t *i2, rc_fv_t *poly )
{
	rc_fv_t fv[2][8];
	int k, pingpong;
	fv[0][0] = *i0; fv[0][1] = *i1; fv[0][2] = *i2;
	unsigned clipflags = fv[0][0].flags | fv[0][1].flags | fv[0][2].flags;
	if( clipflags & RC_ALIAS_Z_CLIP )
	{
		k = rc_alias_clip( vw, fv[0], fv[1], RC_ALIAS_Z_CLIP, 3, 0 );
		if( k == 0 )
			return 0;
		pingpong = 1;
		/* only the first three vertices are looked at, even when the z clip made four (r_aclip.c:239) */
		clipflags = fv[1][0].flags | fv[1][1].flags | fv[1][2].flags;
	}
	else
	{
		pingpong = 0;
		k = 3;
	}
	if( clipflags & RC_ALIAS_LEFT_CLIP )
	{
		k = rc_alias_clip( vw, fv[pingpong], fv[pingpong ^ 1], RC_ALIAS_LEFT_CLIP, k, 1 );
		if( k == 0 ) return 0;
		pingpong ^= 1;
	}
	if( clipflags & RC_ALIAS_RIGHT_CLIP )
	{
		k = rc_alias_clip( vw, fv[pingpong], fv[pingpong ^ 1], RC_ALIAS_RIGHT_CLIP, k, 2 );
		if( k == 0 ) return 0;
		pingpong ^= 1;
	}
	if( clipflags & RC_ALIAS_BOTTOM_CLIP )
	{
		k = rc_alias_clip( vw, fv[pingpong], fv[pingpong ^ 1], RC_ALIAS_BOTTOM_CLIP, k, 3 );
		if( k == 0 ) return 0;
		pingpong ^= 1;
	}
	if( clipflags & RC_ALIAS_TOP_CLIP )
	{
		k = rc_alias_clip( vw, fv[pingpong], fv[pingpong ^ 1], RC_ALIAS_TOP_CLIP, k, 4 );
		if( k == 0 ) return 0;
		pingpong ^= 1;
	}
	const int32_t x0 = vw->vrect[0], y0 = vw->vrect[1], x1 = vw->vrect[0] + vw->vrect[2], y1 = vw->vrect[1] + vw->vrect[3];
	for( int i = 0; i < k; i++ )
	{
		rc_fv_t *p = &fv[pingpong][i];
		if( p->u < x0 ) p->u = x0; else if( p->u > x1 ) p->u = x1;
		if( p->v < y0 ) p->v = y0; else if( p->v > y1 ) p->v = y1;
		p->flags = 0;
		poly[i] = *p;
	}
	return k;
}

/* ------------------------------------------------------------------------------------------ */
/* triangle set-up, r_polyse.c:107-720                                                          */

typedef struct
{
	int32_t ystart, height;      /* first scanline, scanlines */
	int32_t u0;                  /* pixel column at ystart */
	int32_t ubasestep, eup, edown;
	/* left edges only: accumulators at ystart and their per-scanline steps (base step; an "extra" step adds the x gradient) */
	int32_t s0, t0, l0, zi0;
	int32_t sbase, tbase, lbase, zibase, lextra_add;
	int32_t sextra, textra;      /* an "extra" step of s, t: the base step + the x gradient, summed in 32 bits as the reference does */
} rc_triedge_t;

typedef struct
{
	int32_t valid;
	int32_t nleft, nright;
	rc_triedge_t left[2], right[2];
	int32_t right_u_ref;         /* prighttop[0] of the first right edge: spans count from it */
	int32_t sstepx, tstepx, lstepx, zistepx;
	int32_t skinwidth, skinsize; /* row pitch and texel count of the skin (R_PolysetCheckBounds) */
} rc_trisetup_t;

/* floor division for a positive denominator: adivtab (r_polyse.c:91, "adivtab.h") and FloorDivMod (:227) agree with it
 * for every edge of a screen-sized triangle */
RC_HD void rc_floordivmod( int32_t numer, int32_t denom, int32_t *q, int32_t *r )
{
	int32_t qq = numer / denom, rr = numer - qq * denom;
	if( rr < 0 ) { qq--; rr += denom; }
	*q = qq; *r = rr;
}

/* R_PolysetSetUpForLineScan, r_polyse.c:288 */
RC_HD void rc_linescan( int32_t su, int32_t sv, int32_t eu, int32_t ev, rc_triedge_t *e )
{
	rc_floordivmod( eu - su, ev - sv, &e->ubasestep, &e->eup );
	e->edown = ev - sv;
}

/* number of "extra" steps among the first k steps of an edge (errorterm starts at -1, r_polyse.c:296) */
RC_HD int32_t rc_edge_extra( const rc_triedge_t *e, int32_t k )
{
	if( k <= 0 || e->eup <= 0 )
		return 0;
	return (int32_t)(((long long)k * e->eup - 1 ) / e->edown ) + 1;
}

/* R_DrawTriangle + R_PolysetSetEdgeTable + R_PolysetCalcGradients + the set-up half of R_RasterizeAliasPolySmooth.
 * a, b, c: the triangle's vertices in submission order.  Returns 0 when the reference draws nothing. */
RC_HD int rc_tri_setup( const rc_fv_t *a, const rc_fv_t *b, const rc_fv_t *c, int32_t skinwidth, int32_t skinheight, rc_trisetup_t *ts )
{
	ts->valid = 0;
	const int32_t dv0_ab = a->u - b->u, dv1_ab = a->v - b->v;
	if( !( dv0_ab | dv1_ab )) return 0;
	const int32_t dv0_ac = a->u - c->u, dv1_ac = a->v - c->v;
	if( !( dv0_ac | dv1_ac )) return 0;
	const int32_t xdenom = ( dv0_ac * dv1_ab ) - ( dv0_ab * dv1_ac );
	if( !( xdenom < 0 )) return 0;

	int32_t p[3][6];
	const rc_fv_t *src[3] = { a, b, c };
	for( int i = 0; i < 3; i++ )
	{
		p[i][0] = src[i]->u; p[i][1] = src[i]->v; p[i][2] = src[i]->s; p[i][3] = src[i]->t; p[i][4] = src[i]->l; p[i][5] = src[i]->zi;
	}
	/* R_PolysetSetEdgeTable, r_polyse.c:107 + the table at :49-62: {isflattop, numleft, l0, l1, l2, numright, r0, r1, r2} */
	static const signed char tab[12][9] = {
		{ 0, 1, 0, 2, -1, 2, 0, 1, 2 }, { 0, 2, 1, 0, 2, 1, 1, 2, -1 }, { 1, 1, 0, 2, -1, 1, 1, 2, -1 }, { 0, 1, 1, 0, -1, 2, 1, 2, 0 },
		{ 0, 2, 0, 2, 1, 1, 0, 1, -1 }, { 0, 1, 2, 1, -1, 1, 2, 0, -1 }, { 0, 1, 2, 1, -1, 2, 2, 0, 1 }, { 0, 2, 2, 1, 0, 1, 2, 0, -1 },
		{ 0, 1, 1, 0, -1, 1, 1, 2, -1 }, { 1, 1, 2, 1, -1, 1, 0, 1, -1 }, { 1, 1, 1, 0, -1, 1, 2, 0, -1 }, { 0, 1, 0, 2, -1, 1, 0, 1, -1 },
	};
	int idx = 0, done = 0;
	if( p[0][1] >= p[1][1] )
	{
		if( p[0][1] == p[1][1] )
		{
			idx = p[0][1] < p[2][1] ? 2 : 5;
			done = 1;
		}
		else idx = 1;
	}
	if( !done )
	{
		if( p[0][1] == p[2][1] ) { idx = idx ? 8 : 9; done = 1; }
		else if( p[1][1] == p[2][1] ) { idx = idx ? 10 : 11; done = 1; }
	}
	if( !done )
	{
		if( p[0][1] > p[2][1] ) idx += 2;
		if( p[1][1] > p[2][1] ) idx += 4;
	}
	const signed char *et = tab[idx];

	/* R_PolysetCalcGradients, r_polyse.c:325 */
	{
		const float p00_minus_p20 = (float)( p[0][0] - p[2][0] ), p01_minus_p21 = (float)( p[0][1] - p[2][1] );
		const float p10_minus_p20 = (float)( p[1][0] - p[2][0] ), p11_minus_p21 = (float)( p[1][1] - p[2][1] );
		const float xstepdenominv = 1.0f / (float)xdenom, ystepdenominv = -xstepdenominv;
		float t0 = (float)( p[0][4] - p[2][4] ), t1 = (float)( p[1][4] - p[2][4] );
		/* ceil() of the float product promoted to double */
		ts->lstepx = (int32_t)ceil((double)(( t1 * p01_minus_p21 - t0 * p11_minus_p21 ) * xstepdenominv ));
		const int32_t lstepy = (int32_t)ceil((double)(( t1 * p00_minus_p20 - t0 * p10_minus_p20 ) * ystepdenominv ));
		t0 = (float)( p[0][2] - p[2][2] ); t1 = (float)( p[1][2] - p[2][2] );
		ts->sstepx = rc_f2i(( t1 * p01_minus_p21 - t0 * p11_minus_p21 ) * xstepdenominv );
		const int32_t sstepy = rc_f2i(( t1 * p00_minus_p20 - t0 * p10_minus_p20 ) * ystepdenominv );
		t0 = (float)( p[0][3] - p[2][3] ); t1 = (float)( p[1][3] - p[2][3] );
		ts->tstepx = rc_f2i(( t1 * p01_minus_p21 - t0 * p11_minus_p21 ) * xstepdenominv );
		const int32_t tstepy = rc_f2i(( t1 * p00_minus_p20 - t0 * p10_minus_p20 ) * ystepdenominv );
		t0 = (float)( p[0][5] - p[2][5] ); t1 = (float)( p[1][5] - p[2][5] );
		ts->zistepx = rc_f2i(( t1 * p01_minus_p21 - t0 * p11_minus_p21 ) * xstepdenominv );
		const int32_t zistepy = rc_f2i(( t1 * p00_minus_p20 - t0 * p10_minus_p20 ) * ystepdenominv );

		/* left edges, r_polyse.c:391-560 */
		ts->nleft = et[1];
		for( int e = 0; e < ts->nleft; e++ )
		{
			const int32_t *top = p[et[2 + e]], *bot = p[et[3 + e]];
			rc_triedge_t *le = &ts->left[e];
			le->ystart = top[1];
			le->height = bot[1] - top[1];
			le->u0 = top[0];
			/* the second left edge restarts the texel fractions at zero (r_polyse.c:496-497) */
			le->s0 = e ? (int32_t)((uint32_t)top[2] & 0xFFFF0000u ) : top[2];
			le->t0 = e ? (int32_t)((uint32_t)top[3] & 0xFFFF0000u ) : top[3];
			le->l0 = top[4];
			le->zi0 = top[5];
			le->ubasestep = 0; le->eup = 0; le->edown = 1;
			le->sbase = le->tbase = le->lbase = le->zibase = 0; le->lextra_add = 0; le->sextra = le->textra = 0;
			if( le->height != 1 )
			{
				if( le->height <= 0 )
					return 0; /* the reference would walk off its span array; such a table never comes out of a front-facing triangle */
				rc_linescan( top[0], top[1], bot[0], bot[1], le );
				/* for negative steps in x along the left edge, bias toward overflow rather than underflow */
				const int32_t working_lstepx = le->ubasestep < 0 ? ts->lstepx - 1 : ts->lstepx;
				le->sbase = (int32_t)((uint32_t)sstepy + (uint32_t)ts->sstepx * (uint32_t)le->ubasestep );
				le->tbase = (int32_t)((uint32_t)tstepy + (uint32_t)ts->tstepx * (uint32_t)le->ubasestep );
				le->lbase = (int32_t)((uint32_t)lstepy + (uint32_t)working_lstepx * (uint32_t)le->ubasestep );
				le->zibase = (int32_t)((uint32_t)zistepy + (uint32_t)ts->zistepx * (uint32_t)le->ubasestep );
				le->lextra_add = working_lstepx;
				le->sextra = (int32_t)((uint32_t)le->sbase + (uint32_t)ts->sstepx );
				le->textra = (int32_t)((uint32_t)le->tbase + (uint32_t)ts->tstepx );
			}
		}
	}
	/* right edges, r_polyse.c:565-606 */
	ts->nright = et[5];
	ts->right_u_ref = p[et[6]][0];
	for( int e = 0; e < ts->nright; e++ )
	{
		const int32_t *top = p[et[6 + e]], *bot = p[et[7 + e]];
		rc_triedge_t *re = &ts->right[e];
		re->ystart = top[1];
		re->height = bot[1] - top[1];
		re->u0 = top[0];
		if( re->height <= 0 )
			return 0;
		rc_linescan( top[0], top[1], bot[0], bot[1], re );
		re->s0 = re->t0 = re->l0 = re->zi0 = re->sbase = re->tbase = re->lbase = re->zibase = re->lextra_add = re->sextra = re->textra = 0;
	}
	ts->skinwidth = skinwidth;
	ts->skinsize = skinwidth * skinheight;
	ts->valid = 1;
	return 1;
}

/* the scanlines the triangle's span drawer visits: [*y0, *y0 + n) */
RC_HD int32_t rc_tri_rows( const rc_trisetup_t *ts, int32_t *y0 )
{
	*y0 = ts->right[0].ystart;
	return ts->right[0].height + ( ts->nright == 2 ? ts->right[1].height : 0 );
}

typedef struct
{
	int32_t x, count;  /* first pixel, pixels (<= 0: nothing) */
	long long s, t;    /* 16.16 texel accumulators at the first pixel (the pointer + fraction pair of the reference) */
	int32_t l, zi;
} rc_trirow_t;

/* row r (0-based from rc_tri_rows' y0) of the triangle: what spanpackage_t holds for it and the span length the right
 * edge gives it */
RC_HD void rc_tri_row( const rc_trisetup_t *ts, int32_t r, rc_trirow_t *row )
{
	/* left edge part and step number inside it */
	const rc_triedge_t *le = &ts->left[0];
	int32_t k = r;
	if( ts->nleft == 2 && r >= ts->left[0].height )
	{
		le = &ts->left[1];
		k = r - ts->left[0].height;
	}
	const int32_t nx = rc_edge_extra( le, k );
	row->x = le->u0 + k * le->ubasestep + nx;
	row->s = (long long)le->s0 + (long long)( k - nx ) * le->sbase + (long long)nx * le->sextra;
	row->t = (long long)le->t0 + (long long)( k - nx ) * le->tbase + (long long)nx * le->textra;
	row->l = (int32_t)((uint32_t)le->l0 + (uint32_t)k * (uint32_t)le->lbase + (uint32_t)nx * (uint32_t)le->lextra_add );
	(void)0;
	row->zi = (int32_t)((uint32_t)le->zi0 + (uint32_t)k * (uint32_t)le->zibase + (uint32_t)nx * (uint32_t)ts->zistepx );
	/* right edge */
	const rc_triedge_t *re = &ts->right[0];
	int32_t kr = r;
	if( ts->nright == 2 && r >= ts->right[0].height )
	{
		re = &ts->right[1];
		kr = r - ts->right[0].height;
	}
	const int32_t xr = re->u0 + kr * re->ubasestep + rc_edge_extra( re, kr );
	row->count = xr - row->x;
}

/* sbase / tbase are 32-bit sums in the reference (int arithmetic) whose 16.16 halves step a pointer and a fraction:
 * the accumulators above add them as signed 32-bit values, which is what pointer + carry arithmetic amounts to */

/* linear texel index of pixel i of a row (ptex - pskin), and R_PolysetCheckBounds (r_polyse.c:697) for the row */
RC_HD long long rc_tri_texel( const rc_trisetup_t *ts, const rc_trirow_t *row, int32_t i )
{
	const long long s = row->s + (long long)i * ts->sstepx, t = row->t + (long long)i * ts->tstepx;
	return ( s >> 16 ) + ( t >> 16 ) * ts->skinwidth;
}

RC_HD int rc_tri_row_inbounds( const rc_trisetup_t *ts, const rc_trirow_t *row )
{
	const long long a = rc_tri_texel( ts, row, 0 );
	if( a < 0 || a >= ts->skinsize ) return 0;
	if( row->count > 1 )
	{
		const long long b = rc_tri_texel( ts, row, row->count - 1 );
		if( b < 0 || b >= ts->skinsize ) return 0;
	}
	return 1;
}

/* ------------------------------------------------------------------------------------------ */
/* span drawers, r_polyse.c:724-1214, for one pixel                                              */


typedef struct
{
	const uint16_t *colormap; /* [32][8192] */
	const uint16_t *alphamap; /* [3][256][1024] */
	const uint8_t  *addmap, *modmap; /* [65536] */
} rc_blendluts_t;

/* BLEND_COLOR / BLEND_ADD / BLEND_ALPHA, r_local.h:1065-1068, with the truncation to pixel_t of the assignments */
RC_HD uint32_t rc_blend_color( const rc_blendluts_t *L, uint32_t src, uint32_t color )
{
	return ((uint32_t)L->modmap[( src & 0xff00 ) | ( color >> 8 )] << 8 | ( src & color & 0xff ) | (( src & 0xff ) >> 3 )) & 0xFFFFu;
}

RC_HD uint32_t rc_blend_add( const rc_blendluts_t *L, uint32_t src, uint32_t screen )
{
	return ((uint32_t)L->addmap[( src & 0xff00 ) | ( screen >> 8 )] << 8 | ( screen & 0xff ) | ( src & 0xff )) & 0xFFFFu;
}

RC_HD uint32_t rc_blend_alpha_low( const rc_blendluts_t *L, uint32_t alpha, uint32_t src, uint32_t screen )
{
	return ( L->alphamap[( alpha << 18 ) | (( src & 0xff00 ) << 2 ) | ( screen >> 6 )] | ( screen & 0x3f )) & 0xFFFFu;
}

RC_HD uint32_t rc_blend_alpha( const rc_blendluts_t *L, uint32_t alpha, uint32_t src, uint32_t dst )
{
	return alpha > 3 ? rc_blend_alpha_low( L, 7 - 1 - alpha, dst, src ) : rc_blend_alpha_low( L, alpha - 1, src, dst );
}

/*
 * One pixel of a span: src = skin texel, light / zi the row accumulators stepped to the pixel, *pc / *pz the frame.
 * drawer RC_DRAW_*; color / alpha: vid.color / vid.alpha of the triangle's state.
 */
RC_HD void rc_tri_shade( const rc_blendluts_t *L, int drawer, uint32_t color, uint32_t alpha, uint32_t src, int32_t light, int32_t zi,
	uint16_t *pc, int16_t *pz )
{
	const int32_t z16 = zi >> 16;
	if( drawer != RC_DRAW_GLOW && !( z16 >= (int32_t)*pz ))
		return;
	switch( drawer )
	{
	case RC_DRAW_FILL:
		*pc = (uint16_t)( L->colormap[( src >> 3 ) | (((uint32_t)light & 0x1F00u ) << 5 )] | ( src & 7 ));
		*pz = (int16_t)z16;
		break;
	case RC_DRAW_ADDITIVE:
	case RC_DRAW_GLOW:
		*pc = (uint16_t)rc_blend_add( L, rc_blend_color( L, src, color ), *pc );
		break;
	case RC_DRAW_TEXBLEND:
	{
		const uint32_t a = src >> 13;
		const uint32_t temp = rc_blend_color( L, ( src << 3 ) & 0xFFFFu, color );
		if( a == 7 ) *pc = (uint16_t)temp;
		else if( a ) *pc = (uint16_t)rc_blend_alpha( L, a, temp, *pc );
		break;
	}
	default:
	{
		const uint32_t temp = rc_blend_color( L, src, color );
		if( alpha == 7 ) *pc = (uint16_t)temp;
		else if( alpha ) *pc = (uint16_t)rc_blend_alpha( L, alpha, temp, *pc );
		break;
	}
	}
}

/* ------------------------------------------------------------------------------------------ */
/* R_DrawSurfaceDecals (ref/soft/r_surf.c:990-1044) for one texel (tx, ty) of a cache block at `mip`: the face's decals in
 * list order over *px.  Returns 1 when a decal wrote the texel. */
RC_HD int rc_decal_texel( const rc_blendluts_t *L, const uint16_t *texels, const rc_ddecal_t *dl, int32_t ndecals, int32_t mip,
	int32_t tx, int32_t ty, uint32_t *px )
{
	int touched = 0;
	for( int32_t k = 0; k < ndecals; k++ )
	{
		const rc_decalmip_t *dm = &dl[k].mip[mip];
		const int32_t u = tx - dm->x, v = ty - dm->y;
		if( !dm->valid || u < 0 || u >= dm->w || v < 0 || v >= dm->height )
			continue;
		const int32_t sv = ( dm->skip + v ) * dm->tspan / dm->h + dm->t1;
		const uint32_t f = (uint32_t)u * dm->fstep;
		uint32_t src = texels[dl[k].tex_ofs + sv * dl[k].texwidth + dm->s1 + (int32_t)( f >> 16 )];
		uint32_t alpha = 7;
		if( dl[k].transparent )
		{
			alpha &= src >> 13;
			src = ( src << 3 ) & 0xFFFFu;
		}
		if( alpha == 0 )
			continue;
		if( alpha < 7 )
		{
			if( *px == 0x0349u ) /* TRANSPARENT_COLOR */
				continue;
			*px = rc_blend_alpha( L, alpha, src, *px );
		}
		else *px = src;
		touched = 1;
	}
	return touched;
}

/* ------------------------------------------------------------------------------------------ */
/* translucent brush entities: one pixel of D_AlphaSpans16 / D_BlendSpans16 / D_AddSpans16 / D_DrawTurbulent8ZSpan
 * (ref/soft/r_scan.c:698,898,1109,59): texel = what the span's s, t select in the cache block (the texture, for water),
 * izi the span's 1/z accumulator at the pixel */
RC_HD void rc_alpha_pixel( const rc_blendluts_t *L, int mode, uint32_t alpha, int turb, uint32_t texel, int32_t izi, uint16_t *pc, int16_t *pz )
{
	const int32_t z16 = izi >> 16;
	if( !( (int32_t)*pz <= z16 ))
		return;
	if( turb )
	{
		*pc = (uint16_t)( alpha == 7 ? texel : rc_blend_alpha( L, alpha, texel, *pc ));
		return;
	}
	if( texel == 0x0349u ) /* TRANSPARENT_COLOR */
		return;
	if( mode == RC_TB_ALPHATEST ) { *pc = (uint16_t)texel; *pz = (int16_t)z16; }
	else if( mode == RC_TB_ADD ) *pc = (uint16_t)rc_blend_add( L, texel, *pc );
	else *pc = (uint16_t)( alpha != 7 ? rc_blend_alpha( L, alpha, texel, *pc ) : texel );
}

/* ------------------------------------------------------------------------------------------ */
/* the HUD pass: one destination pixel of R_DrawStretchPicImplementation / Draw_Fill (ref/soft/r_draw.c:88-135,192-210) */
RC_HD void rc_2d_pixel( const rc_blendluts_t *L, const rc_cmd2d_t *c, const uint16_t *texels, int32_t v, uint32_t u, uint16_t *dest )
{
	uint32_t src;
	int32_t alpha = c->alpha;
	if( c->kind == RC_2D_PIC )
	{
		const uint32_t sv = (uint32_t)(( c->skip + v ) * ( c->t2 - c->t1 ) / c->h + c->t1 );
		const uint32_t fstep = (uint32_t)((( c->s2 - c->s1 ) << 16 ) / c->w );
		const uint32_t f = u * fstep; /* f += fstep, u times */
		src = texels[(size_t)c->tex_ofs + (size_t)sv * (uint32_t)c->texwidth + (uint32_t)c->s1 + ( f >> 16 )];
		if( c->transparent )
		{
			alpha &= (int32_t)( src >> 13 );
			src = ( src << 3 ) & 0xFFFFu;
		}
		if( alpha == 0 )
			return;
		if( (uint32_t)c->color != 0xFFFFu )
			src = rc_blend_color( L, src, (uint32_t)c->color );
	}
	else
	{
		src = (uint32_t)c->color;
		if( alpha == 0 )
			return;
	}
	if( c->rendermode == 5 ) /* kRenderTransAdd */
		*dest = (uint16_t)rc_blend_add( L, src, *dest );
	else if( c->kind == RC_2D_PIC && c->rendermode == 0x1000 ) /* kRenderScreenFadeModulate */
		*dest = (uint16_t)rc_blend_color( L, *dest, (uint32_t)c->color );
	else if( alpha < 7 )
		*dest = (uint16_t)rc_blend_alpha( L, (uint32_t)alpha, src, *dest );
	else
		*dest = (uint16_t)src;
}

/* ------------------------------------------------------------------------------------------ */
/* studio models: R_StudioDrawPoints' per-vertex and per-normal work (ref/soft/r_studio.c:1802-1895)  */

/* Matrix3x4_VectorTransform (public/xash3d_mathlib.h:255): a studio vertex through its bone */
RC_HD void rc_studio_vertex( const float *m, const float *v, float *out )
{
	out[0] = v[0] * m[0] + v[1] * m[1] + v[2] * m[2] + m[3];
	out[1] = v[0] * m[4] + v[1] * m[5] + v[2] * m[6] + m[7];
	out[2] = v[0] * m[8] + v[1] * m[9] + v[2] * m[10] + m[11];
}

/* TriTexCoord2f (r_triapi.c:180): texture coordinate -> texel of the bound skin */
RC_HD void rc_tri_texcoord( float u, float v, int32_t skinwidth, int32_t skinheight, int32_t *ps, int32_t *pt )
{
	double u1 = 0, v1 = 0;
	u = fmodf( u, 10 );
	v = fmodf( v, 10 );
	if( u < 1000 && u > -1000 ) u1 = u;
	if( v < 1000 && v > -1000 ) v1 = v;
	while( u1 < 0 ) u1 = u1 + 1;
	while( v1 < 0 ) v1 = v1 + 1;
	while( u1 > 1 ) u1 = u1 - 1;
	while( v1 > 1 ) v1 = v1 - 1;
	u1 = u1 < 0.01 ? 0.01 : ( u1 > 0.99 ? 0.99 : u1 );
	v1 = v1 < 0.01 ? 0.01 : ( v1 > 0.99 ? 0.99 : v1 );
	*ps = (int32_t)( skinwidth * u1 );
	*pt = (int32_t)( skinheight * v1 );
}

/* VectorNormalize (public/xash3d_mathlib.h:131): the length through the double-precision sqrt, rounded to float */
RC_HD void rc_vec_normalize( float *v )
{
	float ilength = (float)sqrt( (double)rc_dot3( v, v ));
	if( ilength ) ilength = 1.0f / ilength;
	v[0] *= ilength; v[1] *= ilength; v[2] *= ilength;
}

/* R_StudioSetupChrome (r_studio.c:1169) + the coordinate R_StudioDrawChromeMesh hands to TriTexCoord2f (:1786): the texel a
 * chrome mesh shows at a normal depends on where the viewer is - m: the normal's bone (matrix3x4), vw: the view
 * (chrome_origin = vieworigin, r_studio.c:2318; RI.vright) */
RC_HD void rc_studio_chrome_st( const rc_view_t *vw, const float *m, const float *normal, int32_t skinwidth, int32_t skinheight, int32_t *ps, int32_t *pt )
{
	float tmp[3], up[3], right[3], cup[3], cright[3];
	tmp[0] = -vw->origin[0]; tmp[1] = -vw->origin[1]; tmp[2] = -vw->origin[2];
	tmp[0] += m[3]; tmp[1] += m[7]; tmp[2] += m[11];
	rc_vec_normalize( tmp );
	/* CrossProduct( tmp, RI.vright, chromeupvec ) */
	up[0] = tmp[1] * vw->vright[2] - tmp[2] * vw->vright[1];
	up[1] = tmp[2] * vw->vright[0] - tmp[0] * vw->vright[2];
	up[2] = tmp[0] * vw->vright[1] - tmp[1] * vw->vright[0];
	rc_vec_normalize( up );
	right[0] = up[1] * tmp[2] - up[2] * tmp[1];
	right[1] = up[2] * tmp[0] - up[0] * tmp[2];
	right[2] = up[0] * tmp[1] - up[1] * tmp[0];
	rc_vec_normalize( right );
	/* Matrix3x4_VectorIRotate */
	for( int i = 0; i < 3; i++ )
	{
		cup[i] = up[0] * m[i] + up[1] * m[4 + i] + up[2] * m[8 + i];
		cright[i] = right[0] * m[i] + right[1] * m[4 + i] + right[2] * m[8 + i];
	}
	const float cs = ( rc_dot3( normal, cright ) + 1.0f ) * 32.0f, ct = ( rc_dot3( normal, cup ) + 1.0f ) * 32.0f;
	rc_tri_texcoord( cs * ( 1.0f / (float)skinwidth ), ct * ( 1.0f / (float)skinheight ), skinwidth, skinheight, ps, pt );
}

/* R_StudioLighting (r_studio.c:1391) for one normal, then what R_StudioSetColorBegin / TriColor4ub / _TriColor4f
 * (r_studio.c:1509, r_triapi.c:83-141) make of it: the light level (<< 8) every strip vertex with this normal gets */
RC_HD int32_t rc_studio_light( const rc_spart_t *pt, int flags, const float *normal, const float *blightvec, const uint16_t *lightgamma )
{
	float lv[3];
	if( pt->transadd )
		lv[0] = lv[1] = lv[2] = pt->blend;
	else
	{
		float l;
		if( flags & 0x0004 ) /* STUDIO_NF_FULLBRIGHT */
			l = 1.0f;
		else
		{
			float illum = pt->ambientlight;
			if( flags & 0x0001 ) /* STUDIO_NF_FLATSHADE */
				illum += pt->shadelight * 0.8f;
			else
			{
				float lightcos = rc_dot3( normal, blightvec );
				if( lightcos > 1.0f ) lightcos = 1.0f;
				illum += pt->shadelight;
				const float r = 1.4953241f; /* SHADE_LAMBERT */
				lightcos = ( lightcos + ( r - 1.0f )) / r;
				if( lightcos > 0.0f ) illum -= pt->shadelight * lightcos;
				illum = illum > 0.0f ? illum : 0.0f;
			}
			illum = illum < 255.0f ? illum : 255.0f;
			const uint32_t b = (uint32_t)( illum * 4 );
			l = (float)( b >= 1024 ? 0u : (uint32_t)lightgamma[b] ) / 1023.0f;
		}
		lv[0] = pt->lightcolor[0] * l; lv[1] = pt->lightcolor[1] * l; lv[2] = pt->lightcolor[2] * l;
	}
	float rgb[3];
	for( int i = 0; i < 3; i++ )
	{
		const uint32_t c = (uint32_t)rc_f2i( lv[i] * 255.0f ) & 0xFFu; /* float -> byte */
		rgb[i] = (float)c * ( 1.0f / 255.0f );
	}
	uint32_t light = (uint32_t)rc_f2i(( rgb[0] + rgb[1] + rgb[2] ) * 31 / 3 );
	if( light > 31 ) light = 31;
	return (int32_t)( light << 8 );
}

/* ------------------------------------------------------------------------------------------ */
/* draw items -> world-space corners                                                            */

/* R_DrawSpriteModel's orientation switch + R_DrawSpriteQuad (ref/soft/r_sprite.c:120-266) for one view: the four
 * corners of the quad in submission order; 0 = the sprite is not drawn in this view */
RC_HD int rc_sprite_corners( const rc_view_t *vw, const rc_item_t *it, rc_corner_t *q )
{
	float v_right[3], v_up[3];
	switch( it->sprtype )
	{
	case RC_SPR_ORIENTED:
		for( int i = 0; i < 3; i++ ) { v_right[i] = it->vright[i]; v_up[i] = it->vup[i]; }
		break;
	case RC_SPR_FACING_UPRIGHT:
	{
		v_right[0] = it->origin[1] - vw->origin[1]; v_right[1] = -( it->origin[0] - vw->origin[0] ); v_right[2] = 0.0f;
		v_up[0] = 0.0f; v_up[1] = 0.0f; v_up[2] = 1.0f;
		float ilength = sqrtf( rc_dot3( v_right, v_right ));
		if( ilength ) ilength = 1.0f / ilength;
		v_right[0] *= ilength; v_right[1] *= ilength; v_right[2] *= ilength;
		break;
	}
	case RC_SPR_FWD_PARALLEL_UPRIGHT:
	{
		const float dot = vw->vforward[2];
		if(( dot > 0.999848f ) || ( dot < -0.999848f ))
			return 0;
		v_up[0] = 0.0f; v_up[1] = 0.0f; v_up[2] = 1.0f;
		v_right[0] = vw->vforward[1]; v_right[1] = -vw->vforward[0]; v_right[2] = 0.0f;
		float ilength = sqrtf( rc_dot3( v_right, v_right ));
		if( ilength ) ilength = 1.0f / ilength;
		v_right[0] *= ilength; v_right[1] *= ilength; v_right[2] *= ilength;
		break;
	}
	case RC_SPR_FWD_PARALLEL_ORIENTED:
		for( int i = 0; i < 3; i++ )
		{
			v_right[i] = ( vw->vright[i] * it->cr + vw->vup[i] * it->sr );
			v_up[i] = vw->vright[i] * -it->sr + vw->vup[i] * it->cr;
		}
		break;
	default:
		for( int i = 0; i < 3; i++ ) { v_right[i] = vw->vright[i]; v_up[i] = vw->vup[i]; }
		break;
	}
	const float su[4] = { it->fr_down, it->fr_up, it->fr_up, it->fr_down };
	const float sl[4] = { it->fr_left, it->fr_left, it->fr_right, it->fr_right };
	for( int k = 0; k < 4; k++ )
	{
		float point[3];
		for( int i = 0; i < 3; i++ ) point[i] = it->origin[i] + ( su[k] * it->scale ) * v_up[i];
		for( int i = 0; i < 3; i++ ) point[i] = point[i] + ( sl[k] * it->scale ) * v_right[i];
		q[k].xyz[0] = point[0]; q[k].xyz[1] = point[1]; q[k].xyz[2] = point[2];
		q[k].l = it->light;
		q[k].s = ( k < 2 ) ? it->s0 : it->s1;
		q[k].t = ( k == 0 || k == 3 ) ? it->t1 : it->t0;
	}
	return 1;
}

/* CL_DrawParticles (ref/soft/r_part.c:47): the quad of one particle for a view - it faces the view (cull_vright / cull_vup) and
 * grows with the distance along the view direction so that it does not vanish; corners in TriVertex3f order */
RC_HD int rc_particle_corners( const rc_view_t *vw, const rc_item_t *it, rc_corner_t *q )
{
	const float partsize = it->scale;
	float size = partsize;
	size += ( it->origin[0] - vw->origin[0] ) * vw->vforward[0];
	size += ( it->origin[1] - vw->origin[1] ) * vw->vforward[1];
	size += ( it->origin[2] - vw->origin[2] ) * vw->vforward[2];
	if( size < 20.0f ) size = partsize;
	else size = partsize + size * 0.002f;
	float right[3], up[3];
	for( int k = 0; k < 3; k++ ) { right[k] = vw->vright[k] * size; up[k] = vw->vup[k] * size; }
	for( int k = 0; k < 3; k++ )
	{
		q[0].xyz[k] = it->origin[k] - right[k] + up[k];
		q[1].xyz[k] = it->origin[k] + right[k] + up[k];
		q[2].xyz[k] = it->origin[k] + right[k] - up[k];
		q[3].xyz[k] = it->origin[k] - right[k] - up[k];
	}
	/* TriTexCoord2f( 0, 1 ), ( 0, 0 ), ( 1, 0 ), ( 1, 1 ) */
	q[0].s = it->s0; q[0].t = it->t1; q[1].s = it->s0; q[1].t = it->t0; q[2].s = it->s1; q[2].t = it->t0; q[3].s = it->s1; q[3].t = it->t1;
	for( int k = 0; k < 4; k++ ) q[k].l = it->light;
	return 1;
}

/* R_WorldToScreen (r_main.c:210): x, y of a world point in the view's clip space; not divided for points behind the eye */
RC_HD void rc_world_to_screen( const rc_view_t *vw, const float *p, float *screen )
{
	screen[0] = vw->wvp[0][0] * p[0] + vw->wvp[0][1] * p[1] + vw->wvp[0][2] * p[2] + vw->wvp[0][3];
	screen[1] = vw->wvp[1][0] * p[0] + vw->wvp[1][1] * p[1] + vw->wvp[1][2] * p[2] + vw->wvp[1][3];
	const float w = vw->wvp[2][0] * p[0] + vw->wvp[2][1] * p[1] + vw->wvp[2][2] * p[2] + vw->wvp[2][3];
	screen[2] = 0.0f;
	if( !( w < 0.001f ))
	{
		const float invw = 1.0f / w;
		screen[0] *= invw;
		screen[1] *= invw;
	}
}

/* a tracer's quad for a view (CL_DrawTracers, r_part.c:183-232): the streak start .. start + delta, widened at right
 * angles to its direction on this view's screen.  Corners in the order the reference submits them (TRI_QUADS). */
RC_HD int rc_tracer_corners( const rc_view_t *vw, const rc_item_t *it, rc_corner_t *q )
{
	const float *start = it->origin, *delta = it->angles;
	float end[3], screen[3], screen_last[3], tmp[3], normal[3], tmp2[3], verts[4][3];
	for( int k = 0; k < 3; k++ ) end[k] = start[k] + delta[k];
	rc_world_to_screen( vw, start, screen );
	rc_world_to_screen( vw, end, screen_last );
	tmp[0] = screen[0] - screen_last[0]; tmp[1] = screen[1] - screen_last[1]; tmp[2] = 0.0f;
	rc_vec_normalize( tmp );
	const float su = tmp[0] * it->scale, sr = -tmp[1] * it->scale;
	for( int k = 0; k < 3; k++ )
	{
		normal[k] = vw->vup[k] * su;
		tmp2[k] = vw->vright[k] * sr;
		normal[k] = normal[k] - tmp2[k];
		verts[0][k] = start[k] - normal[k];
		verts[1][k] = start[k] + normal[k];
		verts[2][k] = verts[0][k] + delta[k];
		verts[3][k] = verts[1][k] + delta[k];
	}
	/* TriTexCoord2f( 0, 0.8 ) verts[2], ( 1, 0.8 ) verts[3], ( 1, 0 ) verts[1], ( 0, 0 ) verts[0]; t1 = the texel of 0.8 */
	for( int k = 0; k < 3; k++ ) { q[0].xyz[k] = verts[2][k]; q[1].xyz[k] = verts[3][k]; q[2].xyz[k] = verts[1][k]; q[3].xyz[k] = verts[0][k]; }
	q[0].s = it->s0; q[0].t = it->t1; q[1].s = it->s1; q[1].t = it->t1; q[2].s = it->s1; q[2].t = it->t0; q[3].s = it->s0; q[3].t = it->t0;
	for( int k = 0; k < 4; k++ ) q[k].l = it->light;
	return 1;
}

/* the items that are one quad built per view */
RC_HD int rc_item_is_quad( const rc_item_t *it )
{
	return it->kind == RC_ITEM_SPRITE || it->kind == RC_ITEM_PARTICLE || it->kind == RC_ITEM_TRACER;
}

RC_HD int rc_quad_corners( const rc_view_t *vw, const rc_item_t *it, rc_corner_t *q )
{
	switch( it->kind )
	{
	case RC_ITEM_SPRITE: return rc_sprite_corners( vw, it, q );
	case RC_ITEM_PARTICLE: return rc_particle_corners( vw, it, q );
	default: return rc_tracer_corners( vw, it, q );
	}
}

/*
 * One triangle of a draw item for one view, as the reference draws it: corners -> R_SetupFinalVert ->
 * R_RenderTriangle (trivial reject, R_AliasClipTriangle).  poly receives the vertices of the polygon to draw (3 when
 * nothing was clipped); the triangles are the fan (0, i, i + 1).  Returns the vertex count, 0 = nothing.
 */
RC_HD int rc_view_triangle( const rc_view_t *vw, const rc_corner_t *c0, const rc_corner_t *c1, const rc_corner_t *c2, rc_fv_t *poly )
{
	rc_fv_t a, b, c;
	rc_fv_setup( vw, &a, c0->xyz[0], c0->xyz[1], c0->xyz[2], c0->l, c0->s, c0->t );
	rc_fv_setup( vw, &b, c1->xyz[0], c1->xyz[1], c1->xyz[2], c1->l, c1->s, c1->t );
	rc_fv_setup( vw, &c, c2->xyz[0], c2->xyz[1], c2->xyz[2], c2->l, c2->s, c2->t );
	if( a.flags & b.flags & c.flags )
		return 0; /* completely clipped */
	if( !( a.flags | b.flags | c.flags ))
	{
		poly[0] = a; poly[1] = b; poly[2] = c;
		return 3;
	}
	return rc_alias_clip_triangle( vw, &a, &b, &c, poly );
}

#endif /* RC_TRI_H */
