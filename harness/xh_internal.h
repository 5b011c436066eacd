/* xh_internal.h - shared state of the headless engine stand-in (see xh_host.h) */
#ifndef XH_INTERNAL_H
#define XH_INTERNAL_H

#define ENGINE_DLL 1 /* we are the engine side of the ABI: gives us convar_t (common/cvardef.h:96) */

#include "build.h"
#include "xash3d_types.h"
#include "const.h"
#include "com_model.h"
#include "cl_entity.h"
#include "entity_types.h"
#include "lightstyle.h"
#include "pmove.h"
#include "ref_api.h"
#include "ref_params.h"

#define LM_SAMPLE_SIZE      16 /* engine/common/mod_local.h:24 */
#define LM_SAMPLE_EXTRASIZE 8  /* engine/common/mod_local.h:25 */
#include "xh_host.h"

#define XH_MAX_ENTITIES 4096

typedef struct
{
	void            *dll;
	ref_interface_t ref;
	ref_globals_t   refState;
	ref_client_t    cl;
	ref_host_t      host;
	struct movevars_s movevars;
	render_interface_t drawfuncs;
	dlight_t        dlights[MAX_DLIGHTS];
	dlight_t        elights[MAX_ELIGHTS];
	lightstyle_t    lightstyles[MAX_LIGHTSTYLES];
	color24         palette[256];
	cl_entity_t     viewent;
	cl_entity_t     *entities;
	uint            max_entities;
	int             num_entities;

	byte            texgamma[256];
	uint16_t        lightgamma[1024];
	uint16_t        lineargamma[1024];
	uint16_t        screengamma[1024];

	convar_t        *cvars;
	poolhandle_t    npools;
	uint32_t        rng;
	uint64_t        callbacks;
	int             verbose;

	void            *swbuffer;
	int             sw_w, sw_h;
	unsigned        sw_bpp, sw_mask[3];

	/* world */
	model_t         *world;
	model_t         *submodels; /* array of numsubmodels model_t, [0] == world */
	size_t          visbytes;
	byte            *visscratch;
} xh_state_t;

extern xh_state_t xh;

void *xh_alloc( size_t size );

/* xh_model.c */
int xh_Mod_SampleSizeForFace( const struct msurface_s *surf );
qboolean xh_Mod_BoxVisible( const vec3_t mins, const vec3_t maxs, const byte *visbits );
mleaf_t *xh_Mod_PointInLeaf( const vec3_t p, mnode_t *node, struct model_s *mod );
int xh_R_FatPVS( const float *org, float radius, byte *visbuffer, qboolean merge, qboolean fullvis );
rgbdata_t *xh_FS_LoadImage( const char *filename, const byte *buffer, size_t size );
rgbdata_t *xh_FS_CopyImage( const rgbdata_t *in );
void xh_FS_FreeImage( rgbdata_t *pack );

#endif
