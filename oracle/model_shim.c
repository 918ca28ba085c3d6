/* oracle/model_shim.c -- wraps one generated model header (C dialect) for the CPU oracle.
 * Test infrastructure, not the product path.  Build: see oracle/Makefile. */
#include <math.h>
#include "pbo_model.h"
#define PB_TABLE
#define PB_ALIGN16
#define PB_FN static
#define PB_RESTRICT restrict
#define PB_TID 0
#define PB_TEAM 1
#define PB_WARP 0
#define PB_NWARPS 1
#define PB_LANE 0
#define PB_LANES 1
#define PB_SYNC()
#define PB_LD(x) (x)
#define PB_LDU(x) (x)
#define PB_NAN NAN
#define PB_INF INFINITY
#include PBM_HEADER
static const pbo_model the_model = {
    PBM_N, PBM_NDIFF, PBM_NP_REAL, PBM_NEV, PBM_NNZ, PBM_NTMP, PBM_NVAR,
    pbm_jac_rows, pbm_jac_cols, pbm_res, pbm_jac, pbm_events, pbm_ic, pbm_vars, PBM_NAME};
const pbo_model* pbo_get_model(void) { return &the_model; }
