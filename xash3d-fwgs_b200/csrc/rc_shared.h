/*
 * rc_shared.h - plain-old-data layouts shared by the C host side of ref_cuda (rc_host.c, engine
 * headers, gcc) and the CUDA side (rc_device.cu, nvcc).  No engine types here: the host flattens
 * model_t / msurface_t / mnode_t / mleaf_t (common/com_model.h) into these SoA-friendly records
 * once per R_NewMap and uploads them; kernels never see an engine pointer.
 *
 * Naming follows the reference's domain (ref/soft/r_local.h): surfaces, edges, spans, surface
 * cache blocks, light styles.
 */
#ifndef RC_SHARED_H
#define RC_SHARED_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RC_MAXLIGHTMAPS 4
#define RC_NUM_MIPS     4

/* msurface_t.flags bits the path looks at (common/com_model.h:227-241) */
#define RC_SURF_PLANEBACK   ( 1u << 1 )
#define RC_SURF_DRAWSKY     ( 1u << 2 )
#define RC_SURF_DRAWTURB    ( 1u << 4 )
#define RC_SURF_DRAWTILED   ( 1u << 5 )
#define RC_SURF_CONVEYOR    ( 1u << 6 )
#define RC_SURF_TRANSPARENT ( 1u << 8 )

#define RC_TRANSPARENT_COLOR 0x0349 /* ref/soft/r_local.h:588 */
#define RC_CONTENTS_SOLID    ( -2 )

typedef struct
{
	float normal[3];
	float dist;
	int32_t type; /* 0,1,2 = axial (common/bspfile.h PLANE_X..Z), else general */
	int32_t pad[3];
} rc_plane_t; /* 32 B */

/* children: >= 0 node index, < 0 -> leaf index = -1 - child (BSP30 disk convention, common/bspfile.h:259) */
typedef struct
{
	int32_t plane;
	int32_t child[2];
	int32_t parent; /* -1 for the root */
	int32_t firstsurface, numsurfaces;
	float   minmaxs[6];
} rc_node_t; /* 48 B */

typedef struct
{
	int32_t contents;
	int32_t cluster;
	int32_t parent; /* node index */
	int32_t firstmark, nummark;
	int32_t visofs; /* byte offset into visdata, -1 = no row (all visible) */
	float   minmaxs[6];
} rc_leaf_t; /* 48 B */

typedef struct
{
	int32_t plane;
	uint32_t flags;
	int32_t firstedge, numedges;
	int32_t texnum;      /* index into rc_texture_t table (R_TextureAnimation resolved on host per timestamp) */
	int32_t lightofs;    /* color24 index into lightdata (msurface_t.samples - lightdata), -1 = none */
	int32_t sample_size; /* Mod_SampleSizeForFace */
	uint8_t styles[RC_MAXLIGHTMAPS];
	int16_t texturemins[2], extents[2];
	int16_t lightmapmins[2], lightextents[2];
	float   lmvecs[2][4]; /* mextrasurf_t.lmvecs (texinfo vecs for plain faces) */
	int32_t texflags;     /* mtexinfo_t.flags */
	int32_t cache_ofs[RC_NUM_MIPS]; /* texel offset of the (face,mip) block in the surface cache arena */
	int32_t pad[2];
} rc_surface_t; /* 112 B */

typedef struct
{
	int32_t width, height;
	int32_t ofs[RC_NUM_MIPS]; /* pixel_t offset of each mip in the texel arena */
	int32_t alpha_ofs;        /* alpha plane (TF_HAS_ALPHA) or -1 */
	int32_t flags;
} rc_texture_t; /* 32 B */

/* source of a texture for the device-side encode (k_tex_encode): its RGBA8 image in the batch's source arena */
#define RC_TEXSRC_COLORKEY 1 /* IMAGE_ONEBIT_ALPHA and sw_noalphabrushes off: alpha < 128 becomes TRANSPARENT_COLOR at mip 0 */
typedef struct
{
	uint32_t rgba_ofs;  /* first texel (uint32 RGBA, byte order R G B A) in the source arena */
	int32_t  flags;     /* RC_TEXSRC_* */
} rc_texsrc_t;

/* static world, device pointers (filled by rc_dev_upload_scene) */
typedef struct
{
	int32_t numplanes, numnodes, numleafs, numsurfaces, nummarks, numsurfedges, numedges, numverts;
	int32_t visbytes, visdata_len, lightdata_len, numtextures;
	int32_t numclusters, headnode, maxdepth, pad1; /* maxdepth: BSP tree depth (sizes the walk stack) */
	const rc_plane_t   *planes;
	const rc_node_t    *nodes;
	const rc_leaf_t    *leafs;
	const rc_surface_t *surfaces;
	const int32_t      *marksurfaces;
	const int32_t      *surfedges;
	const uint16_t     *edges;     /* 2 vertex indices per edge */
	const float        *verts;     /* 3 floats per vertex */
	const uint8_t      *visdata;
	const uint8_t      *lightdata; /* color24 */
	const rc_texture_t *textures;
	const uint16_t     *texels;    /* pixel_t arena */
	const uint16_t     *colormap;  /* [32][8192], ref/soft/r_glblit.c:440 */
	const uint16_t     *lightgamma;/* [1024] engine lightgammatable */
	const int32_t      *sintable;  /* [1280] ref/soft/r_main.c:1312 */
	uint16_t           *cache;     /* surface cache arena */
	/* derived on upload (rc_dev_upload_scene); host callers leave them NULL */
	const uint32_t     *pvsdense;  /* [numleafs][visbytes_padded / 4] Mod_DecompressPVS of every leaf's row */
	const int32_t      *surfnode;  /* [numsurfaces] the node a surface hangs off */
	const uint32_t     *surfback;  /* [numsurfaces bits] SURF_PLANEBACK */
	const float        *sedgeverts;/* [numsurfedges][6] both vertices of every surfedge in its winding direction: one load
	                                * instead of the surfedges -> edges -> verts chain of three dependent ones */
	const int32_t      *sedgepartner;/* [numsurfedges] the other face that lists the same medge (R_RenderFace's edge cache,
	                                * ref/soft/r_rast.c:447-478, couples the two), bit 30 set: it walks the edge in the same
	                                * direction; -1: none */
} rc_scene_t;

/* an active dynamic light of the batch (dlight_t, common/q_client.h:57, after R_PushDlights' die / radius test) */
#define RC_MAX_DLIGHTS 32
typedef struct
{
	float   origin[3];
	float   radius;
	float   minlight;
	int32_t monolight; /* LightToTexGamma(( r + g + b ) / 3 * 4 ) * 3, ref/soft/r_surf.c:138 */
	int32_t bit;       /* index in the engine's table = bit in msurface_t.dlightbits */
	float   virtual_radius; /* radius * max( 1, r_dlight_virtual_radius ), ref/common/ref_light.c:96 */
} rc_dlight_t; /* 32 B */

/* per-view constants: everything R_SetupFrustum/R_SetupFrameQ/R_ViewChanged/R_TransformFrustum derive
 * (ref/soft/r_main.c:387, r_misc.c:101-322).  Computed on the host with the host libm and the reference's
 * float expression order, so the device never re-derives a transcendental. */
typedef struct
{
	float   origin[3];
	float   vright[3], vup[3], vforward[3];
	float   xscale, yscale, xcenter, ycenter, xscaleinv, yscaleinv;
	float   fvrectx_adj, fvrecty_adj, fvrectright_adj, fvrectbottom_adj;
	int32_t vrect_x_adj_shift20, vrectright_adj_shift20;
	int32_t vrect[4];            /* x, y, width, height */
	float   clip_normal[4][3];   /* view_clipplanes (world space) */
	float   clip_dist[4];
	int32_t frustum_idx[4][6];   /* reject / accept corner selectors */
	float   scale_for_mip, scalemip[3];
	int32_t minmip;
	float   modelorg_t[3];       /* transformed_modelorg: TransformVector( vieworigin ) */
	int32_t turb_phase;          /* (int)(time * SPEED) & (CYCLE - 1) */
	int32_t sky_color;           /* (int)sw_clearcolor & 0xFFFF */
	int32_t flags;               /* ref_viewpass_t.flags */
	float   pvs_radius;
	float   screenedge[4][3];    /* qfrustum.screenedge normals (view space): R_TransformFrustum runs again per brush entity */
	float   wvp[3][4];           /* rows 0, 1 and 3 of RI.worldviewProjectionMatrix (r_main.c:1078-1081): R_WorldToScreen */
	int32_t texfilt;             /* sw_texfilt == 1: D_DrawSpans16's texture dither (ref/soft/r_scan.c:503-672) */
	float   aliastransform[3][4]; /* R_SetUpWorldTransform (ref/soft/r_trialias.c:52): world -> view space for R_SetupFinalVert */
} rc_view_t;

/*
 * Entities (everything that is not a world face goes through the triangle rasteriser, ref/soft/r_polyse.c).  The host
 * turns the frame's entity lists into an ordered list of draw items - the order R_DrawEntitiesOnList (r_main.c:504)
 * draws in; translucent entities come with a sort record, their order is per view (R_TransEntityCompare, r_main.c:157).
 */
enum { RC_ITEM_TRIS = 0, RC_ITEM_SPRITE = 1, RC_ITEM_MESH = 2, RC_ITEM_BRUSH = 3 /* a translucent brush entity: first = its rc_bent_t */,
	RC_ITEM_PARTICLE = 4 /* a particle of CL_DrawParticles: a quad that faces every view (origin, scale = partsize) */,
	RC_ITEM_TRACER = 5 /* a tracer of CL_DrawTracers: a quad along the streak, as wide as its direction on the screen of every
	                    * view says (origin = start, angles = the streak's vector, scale = gTracerSize of its type) */ };
/* the span drawer a texture bind selects (GL_Bind, ref/soft/r_image.c:43-75): R_PolysetFillSpans8, ...Additive, ...Glow,
 * ...TextureBlended, ...Blended (r_polyse.c:724-1214) */
enum { RC_DRAW_FILL = 0, RC_DRAW_ADDITIVE = 1, RC_DRAW_GLOW = 2, RC_DRAW_TEXBLEND = 3, RC_DRAW_BLEND = 4 };
enum { RC_SPR_FWD_PARALLEL_UPRIGHT = 0, RC_SPR_FACING_UPRIGHT = 1, RC_SPR_FWD_PARALLEL = 2, RC_SPR_ORIENTED = 3, RC_SPR_FWD_PARALLEL_ORIENTED = 4 }; /* common/sprite.h */

/* a triangle corner in world space: what TriVertex3f hands to R_SetupFinalVert (r_triapi.c:221) */
typedef struct
{
	float   xyz[3];
	int32_t l;     /* light << 8 */
	int16_t s, t;  /* texel, TriTexCoord2f's result */
} rc_corner_t; /* 20 B */

typedef struct
{
	int32_t kind;                 /* RC_ITEM_* */
	int32_t first, count;         /* TRIS: first corner (3 per triangle) and triangle count; MESH: first static corner, triangle count */
	int32_t drawer;               /* RC_DRAW_* the texture bind selected (GL_Bind, r_image.c:43) */
	int32_t tex_ofs, skinwidth, skinheight; /* r_affinetridesc: texel arena offset of pskin, its size */
	int32_t color, alpha;         /* vid.color, vid.alpha */
	int32_t sortent;              /* translucent entity the item belongs to (index into the sort records), -1: fixed order */
	int32_t order0;               /* number of triangles in the items before this one */
	/* RC_ITEM_SPRITE (R_DrawSpriteModel / R_DrawSpriteQuad, r_sprite.c:120-266) */
	int32_t sprtype, light;
	float   origin[3], scale, angles[3];
	float   fr_up, fr_down, fr_left, fr_right;
	int16_t s0, s1, t0, t1;       /* texels of texture coordinates 0 and 1 */
	float   vright[3], vup[3];    /* SPR_ORIENTED: AngleVectors of the entity's angles (origin already pulled back by 0.01 forward) */
	float   sr, cr;               /* SPR_FWD_PARALLEL_ORIENTED: SinCos of the roll */
	/* RC_ITEM_MESH: the entity's slices of the batch's transformed vertices and per-normal lights */
	int32_t vert_base, light_base;
	int32_t part;                 /* RC_ITEM_MESH: index of its (entity, submodel) in the batch's part list; else -1 */
	int32_t chrome;               /* RC_ITEM_MESH with STUDIO_NF_CHROME: texture coordinates per view from the normals (R_StudioSetupChrome) */
	int32_t glow;                 /* RC_ITEM_SPRITE drawn with kRenderGlow: its column in the batch's per-view glow records; else -1 */
} rc_item_t; /* 164 B */

/* a decal of a face, as R_DrawSurfaceDecals (ref/soft/r_surf.c:914) blits it into a cache block of the face: per mip level
 * the rectangle of the block it covers and the part of the picture that lands there (the reference's integer set-up,
 * done on the host: it depends on the face, the decal and the mip only) */
typedef struct
{
	int32_t  valid;
	int32_t  x, y, w;        /* first column / row and width of the rectangle in the block */
	int32_t  h;              /* the row divisor: sv = ( skip + v ) * tspan / h + t1 */
	int32_t  height, skip;   /* rows drawn, rows of the rectangle above the block */
	int32_t  s1, t1, tspan;  /* first picture column, first picture row, t2 - t1 */
	uint32_t fstep;          /* 16.16 picture columns per block column */
} rc_decalmip_t;

typedef struct
{
	int32_t tex_ofs;         /* texel arena offset of the picture (its alpha plane when it has one) */
	int32_t texwidth;
	int32_t transparent;     /* the picture has an alpha plane: 3-bit alpha in the top bits of a texel */
	rc_decalmip_t mip[4];
} rc_ddecal_t;

/* a kRenderGlow sprite for one view (R_SpriteOccluded / R_SpriteGlowBlend, ref/soft/r_sprite.c:59-117): whether it is drawn
 * (scissor, the engine's occlusion trace from this eye, fade) and the size and colour its distance from this eye gives it.
 * Computed on the host per (view, sprite) - the trace is an engine callback - and laid out [view][sprite]. */
typedef struct
{
	int32_t visible;
	float   scale;
	int32_t light, color, alpha;  /* what _TriColor4f leaves for tr.blend * brightness */
} rc_glow_t; /* 20 B */

/* a studio submodel as the device keeps it (static per model file): R_StudioDrawPoints' inputs (ref/soft/r_studio.c:1802) */
typedef struct
{
	uint16_t vert, norm; /* indices into the submodel's vertices / the submodel's normals (tricmds, r_studio.c:1685) */
	int16_t  s, t;       /* TriTexCoord2f's texel for this strip vertex on the mesh's skin */
} rc_mcorner_t; /* 8 B */

/* one drawn (entity, submodel) of a batch: what k_studio_points needs to transform its vertices and light its normals */
typedef struct
{
	int32_t vert_base, numverts;   /* slice of the batch's transformed vertices */
	int32_t light_base, numnorms;  /* slice of the batch's per-normal lights */
	int32_t sverts, svertbone;     /* static arena offsets (floats / bytes): studio vertices, their bones */
	int32_t snorms, snormbone, snormflags; /* studio normals, their bones, the face flags of the mesh that owns each normal */
	int32_t bones;                 /* first matrix3x4 of the entity in the batch's bone array (12 floats each) */
	int32_t blightvec;             /* first vec3 of the entity's bone-space light vectors */
	float   ambientlight, shadelight, lightcolor[3];
	float   blend;                 /* tr.blend */
	int32_t transadd;              /* the entity draws in kRenderTransAdd: every light value is tr.blend */
	int32_t pad[2];
} rc_spart_t; /* 80 B */

/* a translucent entity: R_TransEntityCompare's inputs (r_main.c:157-200) */
typedef struct
{
	float   center[3];    /* origin + average of the model's mins / maxs */
	int32_t use_dist;     /* 0: distance is 1e9 (a brush model in kRenderTransAlpha) */
	int32_t rank;         /* R_RankForRenderMode */
	int32_t first_item, nitems; /* the entity's draw items (contiguous in the list) */
	int32_t pad[1];
} rc_sortent_t; /* 32 B */

/* an opaque brush entity of the batch (tr.draw_list->edge_entities, R_DrawBEntitiesOnList ref/soft/r_main.c:781): its
 * faces join the view's edge list - keyed by the world leaf they lie in - and are drawn by the world's span stages */
#define RC_MAX_BENTS 256
typedef struct
{
	float   origin[3];       /* r_entorigin */
	float   rot[3][3];       /* entity_rotation (R_RotateBmodel, r_bsp.c:97): yaw, pitch, roll matrices concatenated on the host */
	float   minmaxs[6];      /* RotatedBBox + origin */
	int32_t firstface, numfaces;
	int32_t trans;           /* 0: opaque (edge_entities).  RC_TB_*: a translucent entity, drawn in its turn among the sorted
	                          * entities by R_DrawBrushModel (r_main.c:861) through D_AlphaSurf's span drawers (r_edge.c:904) */
	int32_t alpha;           /* D_AlphaSurf: renderamt * 7 / 255, at least 1 */
} rc_bent_t; /* 96 B */
enum { RC_TB_ALPHATEST = 1 /* kRenderTransAlpha: D_AlphaSpans16 */, RC_TB_ADD = 2 /* kRenderTransAdd: D_AddSpans16 */, RC_TB_BLEND = 3 /* D_BlendSpans16 */ };
#define RC_TB_MAXSURFS 64   /* pieces one translucent brush entity can have in a view */
#define RC_TB_MAXEDGES 384
#define RC_TB_LINESPANS 24  /* spans of one entity on one scanline */

/* a 2-D draw of the frame's HUD pass (R_DrawStretchPic / Draw_Fill, ref/soft/r_draw.c:38,159), clipped on the host */
enum { RC_2D_PIC = 0, RC_2D_FILL = 1 };
typedef struct
{
	int32_t kind;             /* RC_2D_* */
	int32_t x, y, w, h;       /* destination after clipping (h: the unclipped height the row step divides by) */
	int32_t s1, t1, s2, t2;   /* source rectangle in texels */
	int32_t skip, height;     /* rows cut off above the screen, rows drawn */
	int32_t tex_ofs, texwidth;/* pixels[0] or alpha_pixels in the texel arena */
	int32_t transparent;      /* source carries 3 alpha bits on top (alpha_pixels) */
	int32_t color, alpha, rendermode; /* vid.color, vid.alpha, vid.rendermode when the call was made */
	int32_t pad;
} rc_cmd2d_t; /* 72 B */

/* a face the BSP walk decided to draw, in front-to-back order (R_RecursiveWorldNode -> R_RenderFace) */
typedef struct
{
	int32_t surf;      /* index into rc_scene_t.surfaces */
	int32_t clipflags; /* frustum planes still to clip against */
} rc_cand_t;

/* surf_t equivalent (ref/soft/r_local.h:805-824) + the per-surface draw state D_SolidSurf/D_CalcGradients
 * derive (ref/soft/r_edge.c:741-813,983-1053); one per candidate, valid when .emitted != 0 */
typedef struct
{
	int32_t  surf;
	uint32_t flags;
	int32_t  emitted;
	int32_t  miplevel;
	float    nearzi;
	float    d_zistepu, d_zistepv, d_ziorigin;
	float    d_sdivzstepu, d_tdivzstepu, d_sdivzstepv, d_tdivzstepv, d_sdivzorigin, d_tdivzorigin;
	int32_t  sadjust, tadjust, bbextents, bbextentt;
	int32_t  cache_ofs;  /* texel offset of the block (or of the mip-0 texture for turbulent faces) */
	int32_t  cachewidth;
} rc_surf_t; /* 80 B */

/* edge_t equivalent (ref/soft/r_local.h:827-836): u at scanline v, stepped by u_step per line up to v2 */
typedef struct
{
	int32_t  u, u_step;
	uint32_t vv;    /* v | v2 << 16 */
	uint32_t surfs; /* trailing (surfs[0]) | leading (surfs[1]) << 16, candidate index + 1, 0 = none */
} rc_edge_t; /* 16 B */

/* per-batch device work buffers and outputs */
typedef struct
{
	int32_t nviews, width, height, pad;
	int32_t max_cands, max_edges; /* per-view capacities */
	const rc_view_t *views;
	uint8_t   *leafvis;   /* [nviews][visbytes_padded] fat PVS rows (R_MarkLeaves input) */
	uint32_t  *nodevis;   /* [nviews][nodewords] node visframe bits */
	uint32_t  *leafmark;  /* [nviews][leafwords] leaf visframe bits */
	uint32_t  *surfmark;  /* [nviews][surfwords] msurface_t.visframe bits */
	int32_t   *walkstack; /* [nviews][stack_ints] explicit recursion stack */
	rc_cand_t *cands;     /* [nviews][max_cands] */
	rc_surf_t *surfs;     /* [nviews][max_cands] */
	rc_edge_t *edges;     /* [nviews][max_edges] */
	int32_t   *ncands;    /* [nviews] */
	int32_t   *nedges;    /* [nviews] */
	uint32_t  *cache_need;/* [numsurfaces*4 bits] (face,mip) blocks this batch samples */
	uint16_t  *color;     /* [nviews][height][width] pixel_t */
	int16_t   *zbuf;      /* [nviews][height][width] */
	uint64_t  *counters;  /* RC_CTR_* */
	int32_t   visbytes_padded, nodewords, leafwords, surfwords, stack_ints, pad2;
} rc_batch_t;

enum
{
	RC_CTR_VIEWS = 0, RC_CTR_PIXELS, RC_CTR_CANDS, RC_CTR_SURFS, RC_CTR_EDGES, RC_CTR_SPANS,
	RC_CTR_CACHE_BLOCKS, RC_CTR_CACHE_TEXELS, RC_CTR_OVERFLOW_EDGES, RC_CTR_OVERFLOW_EVENTS,
	RC_CTR_OVERFLOW_STACK, RC_CTR_OVERFLOW_CANDS, RC_CTR_SERIAL_LINES, RC_CTR_SORTED_LINES, RC_CTR_UNSTAGED_BANDS,
	RC_CTR_MAX_CANDS /* most faces one view listed */, RC_CTR_MAX_EDGES /* most edges one view emitted */, RC_CTR_MAX_BAND_EVENTS /* most event slots one band wanted */,
	RC_CTR_OVERFLOW_CACHE /* (face, mip) blocks k_cache_build could not build */, RC_CTR_TRIS, RC_CTR_TRI_PIXELS, RC_CTR_WIDE_LINES /* scanlines k_lines_wide took */, RC_CTR_OVERFLOW_FIX, RC_CTR_FIXED_FACES,
	RC_CTR_TRI_WRITTEN /* pixels k_tri_resolve shaded (the key path's visible fragments) */, RC_CTR_BIG_TRIS /* triangles k_tris_fast_big took */,
	RC_CTR_BMODEL_SURFS /* brush entity surfaces (face fragments) posted */, RC_CTR_BMODEL_TIES /* scanlines where two of them with one leaf key overlapped */,
	RC_CTR_OVERFLOW_BMODEL /* a face whose BSP clipping outgrew the fragment scratch */, RC_CTR_COUNT = 64 /* 16.. = debug-build phase clocks */
};

#ifdef __cplusplus
}
#endif
#endif
