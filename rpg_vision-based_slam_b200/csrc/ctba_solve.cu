// placeholder — replaced by the LM solver
#include "ctba_internal.h"
namespace ctba {
struct SolverData {};
void solver_free(SolverData* s) { delete s; }
int solver_num_unknowns(ctba_problem* p) { return 0; }
int solver_solve(ctba_problem* p, int, ctba_summary*) { p->err = "solver not built"; return CTBA_E_STATE; }
}
