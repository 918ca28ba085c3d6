/* oracle/pbo_model.h -- interface between the CPU oracle (test infrastructure) and a
 * model compiled from the generated C dialect (pybamm_b200/codegen.py).
 * Plays the role of the CasADi function set of the reference
 * (idaklu_source/Expressions/Casadi/CasadiFunctions.hpp). */
#ifndef PBO_MODEL_H
#define PBO_MODEL_H
typedef void (*pbo_fn)(const double* y, const double* p, double t, double cj, double* out, double* tmp);
typedef struct {
  int n, n_diff, n_p, n_ev, nnz, n_tmp, n_var;
  const int* jac_rows;
  const int* jac_cols;
  pbo_fn res, jac, events, ic, vars;
  const char* name;
} pbo_model;
const pbo_model* pbo_get_model(void);
#endif
