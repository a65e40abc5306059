/* Serial stand-in for Intel's libiomp5.so — TEST INFRASTRUCTURE ONLY.
 *
 * The reference ships one runnable build of itself, MonteCarlo/MonteCarloOpenMP
 * (x86-64 ELF, ifort 16 runtime linked statically).  Its only unresolved
 * dependency is libiomp5.so, from which it imports the 15 entry points below
 * (nm -D --undefined-only).  The classical LJ path has no OpenMP region at all
 * (reference MonteCarlo/MonteCarlo.F90:792-828 vs the only worksharing loop at
 * :857-859, which is the QM path), so running every "parallel" construct on the
 * calling thread changes nothing for the path this repository rebuilds.
 *
 * Build: see oracle/Makefile (target _ref/libiomp5.so).
 */
#include <stdarg.h>

#define NARGS 24
typedef void (*microtask_t)(int *gtid, int *btid, ...);

void __kmpc_begin(void *loc, int flags) { (void)loc; (void)flags; }
void __kmpc_end(void *loc) { (void)loc; }
int  __kmpc_global_thread_num(void *loc) { (void)loc; return 0; }
int  __kmpc_ok_to_fork(void *loc) { (void)loc; return 1; }
void __kmpc_serialized_parallel(void *loc, int gtid) { (void)loc; (void)gtid; }
void __kmpc_end_serialized_parallel(void *loc, int gtid) { (void)loc; (void)gtid; }
void __kmpc_barrier(void *loc, int gtid) { (void)loc; (void)gtid; }
void __kmpc_critical(void *loc, int gtid, void *crit) { (void)loc; (void)gtid; (void)crit; }
void __kmpc_end_critical(void *loc, int gtid, void *crit) { (void)loc; (void)gtid; (void)crit; }

/* dynamic/runtime schedule: hand the whole iteration range out exactly once */
static int disp_lb, disp_ub, disp_st, disp_pending;
void __kmpc_dispatch_init_4(void *loc, int gtid, int sched, int lb, int ub, int st, int chunk)
{
    (void)loc; (void)gtid; (void)sched; (void)chunk;
    disp_lb = lb; disp_ub = ub; disp_st = st; disp_pending = 1;
}
int __kmpc_dispatch_next_4(void *loc, int gtid, int *p_last, int *p_lb, int *p_ub, int *p_st)
{
    (void)loc; (void)gtid;
    if (!disp_pending) return 0;
    disp_pending = 0;
    *p_lb = disp_lb; *p_ub = disp_ub; *p_st = disp_st;
    if (p_last) *p_last = 1;
    return 1;
}

/* fork: run the outlined body once, on this thread, as thread 0 of a team of 1 */
void __kmpc_fork_call(void *loc, int argc, microtask_t fn, ...)
{
    void *a[NARGS] = {0};
    va_list ap;
    int gtid = 0, btid = 0, i;
    (void)loc;
    va_start(ap, fn);
    for (i = 0; i < argc && i < NARGS; ++i) a[i] = va_arg(ap, void *);
    va_end(ap);
    fn(&gtid, &btid, a[0], a[1], a[2], a[3], a[4], a[5], a[6], a[7], a[8], a[9], a[10], a[11],
       a[12], a[13], a[14], a[15], a[16], a[17], a[18], a[19], a[20], a[21], a[22], a[23]);
}

int omp_get_num_procs(void) { return 1; }
int omp_get_num_threads(void) { return 1; }
int omp_get_thread_num(void) { return 0; }
