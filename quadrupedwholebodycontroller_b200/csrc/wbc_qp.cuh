// Batched dense active-set QP solver mirroring qpOASES 3.2.2 SQProblem::init / hotstart as configured by the
// reference (Options::setToReliable, nWSR = 100, no simple bounds; src/whole_body_controller.cpp:517-541).
// One QP instance per warp; Q (42x42), R, T and every vector live in shared memory.
//
// What is mirrored (reference lib/qpOASES/src): the online active-set homotopy of
// QProblem::solveQP with far bounds (QProblem.cpp:446-640, 5498-5552), ramping (:5416-5492), drift correction
// (:5559-5652), per-iteration projected Cholesky (:2036-2129), determineStepDirection with <= 2 refinement passes
// (:4637-4975), the six sequential ratio tests (:4981-5278, QProblemB.cpp:2065-2107), add/remove of bounds and
// constraints with Givens maintenance of the TQ factorisation (:2819-4252), full LI tests and LI exchange
// (:2982-3300, 3487-3780), insertion-ordered index lists (Indexlist.cpp), and SQProblem::setupNewAuxiliaryQP
// (SQProblem.cpp:340-515) for the warm start.
//
// What is NOT executed, with the reason it cannot change a result for this QP family (H = I + PSD, finite A):
//  * Givens updates of R inside add/remove: enableCholeskyRefactorisation = 1 recomputes R from Q right after every
//    working-set change and nothing reads R in between except the rho^2 test below  -> dead stores.
//  * flipping bounds (Flipper snapshot of R,Q,T): fires only if rho^2 = z'Hz - r'r <= 1e3*EPS; z is a unit vector and
//    H >= I, so rho^2 >= 1.
//  * zero-curvature (NZC) test: fires only if max|dy| < 3e3*EPS*max|dx| for the removal direction; there
//    dx'H dx = |dy_k| >= |dx|^2 and |dx|_inf >= 1/(42 max|A|), so the test never passes; the step direction it would
//    compute has no side effect.
//  * Hessian regularisation, HST_ZERO / HST_IDENTITY / semidefinite branches: H is always typed HST_POSDEF.
// If one of these preconditions were ever violated the solver reports QP_RET_UNSUPPORTED_PATH instead of guessing.
#pragma once
#include "wbc_common.cuh"
#include <cstddef>
#if !defined(__CUDACC__) && (defined(WBC_QP_TRACE) || defined(WBC_LI_STATS) || defined(WBC_QP_DBG))
#include <cstdio>
#include <cstdlib>
#endif

namespace wbc {
#if !defined(__CUDACC__) && defined(WBC_QP_DBG)      // host emulation only: dump of the step-direction intermediates of one iteration (tools/host_emul)
static int g_dbg_iter=-1;
static int dbg_target(){ static int t=-2; if(t==-2){ const char* e=getenv("DBG_ITER"); t=e?atoi(e):-1; } return t; }
#define WBC_DBG_VEC(name,expr,len) if(g_dbg_iter==dbg_target()){ printf("REF %s",name); for(int i_=0;i_<(len);i_++){ const int i=i_; printf(" %a",(double)(expr)); } printf("\n"); }
#else
#define WBC_DBG_VEC(name,expr,len) do{}while(0)
#endif

// qpOASES returnValue numbering (MessageHandling.hpp) for the codes this path can produce
enum : int { QP_OK=0, QP_RET_DIV_BY_ZERO=1, QP_RET_INIT_FAILED=33, QP_RET_INIT_FAILED_CHOLESKY=35, QP_RET_INIT_FAILED_HOTSTART=36,
             QP_RET_INIT_FAILED_INFEASIBILITY=37, QP_RET_INIT_FAILED_UNBOUNDEDNESS=38, QP_RET_SETUP_AUXILIARYQP_FAILED=43,
             QP_RET_HOTSTART_FAILED_AS_QP_NOT_INITIALISED=54, QP_RET_STEPDIRECTION_DETERMINATION_FAILED=57, QP_RET_HOMOTOPY_STEP_FAILED=60,
             QP_RET_HOTSTART_STOPPED_INFEASIBILITY=61, QP_RET_HOTSTART_STOPPED_UNBOUNDEDNESS=62, QP_RET_MAX_NWSR_REACHED=64, QP_RET_STEPDIRECTION_FAILED_TQ=68, QP_RET_STEPDIRECTION_FAILED_CHOLESKY=69,
             QP_RET_HESSIAN_NOT_SPD=99, QP_RET_UNSUPPORTED_PATH=1001,
             QP_SUSPENDED=2001 };   // not a qpOASES code, never published: the solve was parked at the top of an iteration (cold batches, "parking" below) and continues later, bit for bit
enum : int { QPS_NOTINITIALISED=0, QPS_PREPARINGAUXILIARYQP=1, QPS_AUXILIARYQPSOLVED=2, QPS_PERFORMINGHOMOTOPY=3, QPS_HOMOTOPYQPSOLVED=4, QPS_SOLVED=5 };
enum : int { ST_LOWER=-1, ST_INACTIVE=0, ST_UPPER=1 };     // Types.hpp:272-280

// Constants.hpp:48-64 and Options::setToReliable (Options.cpp:91-189)
#if defined(WBC_FP32)
// qpOASES built with __USE_SINGLE_PRECISION__: Constants.hpp:46-47, Options.cpp:111-152 (setToDefault) and :203-207 (setToReliable: 3 refinement steps)
constexpr real QEPS=real(1.193e-07), QZERO=real(1.0e-25), QINFTY=real(1.0e20);
constexpr real OPT_terminationTolerance=real(1.0e2)*QEPS, OPT_boundTolerance=real(1.0e2)*QEPS, OPT_boundRelaxation=real(1.0e4),
                 OPT_epsNum=-real(1.0e2)*QEPS, OPT_epsDen=real(1.0e2)*QEPS, OPT_maxPrimalJump=real(1.0e8), OPT_maxDualJump=real(1.0e8),
                 OPT_initialFarBounds=real(1.0e6), OPT_growFarBounds=real(1.0e3), OPT_epsFlipping=real(5.0e1)*QEPS, OPT_epsIterRef=real(1.0e2)*QEPS,
                 OPT_epsLITests=real(5.0e1)*QEPS, OPT_ramp0=real(0.5), OPT_ramp1=real(1.0);
constexpr int OPT_numRefinementSteps=3;
// the shortcut of the full linear-independence test only trusts clear cases in single precision: everything below 1e-2 |a| runs the reference's test
constexpr real LI_SURE=real(1.0e-2), LI_NONE=real(-1.0);
#else
constexpr real QEPS=real(2.221e-16), QZERO=real(1.0e-25), QINFTY=real(1.0e20);
constexpr real OPT_terminationTolerance=real(5.0e6)*QEPS, OPT_boundTolerance=real(1.0e6)*QEPS, OPT_boundRelaxation=real(1.0e4),
                 OPT_epsNum=-real(1.0e3)*QEPS, OPT_epsDen=real(1.0e3)*QEPS, OPT_maxPrimalJump=real(1.0e8), OPT_maxDualJump=real(1.0e8),
                 OPT_initialFarBounds=real(1.0e6), OPT_growFarBounds=real(1.0e3), OPT_epsFlipping=real(1.0e3)*QEPS, OPT_epsIterRef=real(1.0e2)*QEPS,
                 OPT_epsLITests=real(1.0e5)*QEPS, OPT_ramp0=real(0.5), OPT_ramp1=real(1.0);
constexpr int OPT_numRefinementSteps=2;
// verdict of the full linear-independence test without running it: max|Z'a| > LI_SURE |a| -> independent, <= LI_NONE |a| -> dependent
constexpr real LI_SURE=real(1.0e-6), LI_NONE=real(1.0e-12);
#endif

#ifndef WBC_PARK_MIN_ITER
  #define WBC_PARK_MIN_ITER 80     /* measured: 0 / 20 / 40 / 50 / 60 / 70 / 80 / 90 / 100 -> 27.7 / 26.7 / 25.2 / 25.1 / 25.0 / 24.8 / 24.3 / 25.5 / 26.3 ms on cold4096 (profiles/r02_gang_variants.txt) */
#endif
#ifndef WBC_PARK_MIN_DONE
  #define WBC_PARK_MIN_DONE 1
#endif
// A cold solve is parked at the top of iteration >= PARK_MIN_ITER once >= PARK_MIN_DONE members of its gang are done.  Cold solves of this QP family take 62..100
// iterations (mean 81): up to iteration 80 the gang stays together even if a member or two idle (lock-step iterations are the cheapest way to run them), after it the
// stragglers are parked and finished by the second launch.  A batch whose solves are all shorter never parks and runs exactly like kernel 2.
constexpr int PARK_MIN_ITER=WBC_PARK_MIN_ITER, PARK_MIN_DONE=WBC_PARK_MIN_DONE;
constexpr int HZLD=42;                // row length of the (HZ)_force block (the 4-wide tiles of zhz_tiles may read up to 2 elements past a row's end and discard them)
constexpr int LD=43;                  // leading dimension of Q (odd: conflict-free row AND column sweeps)
constexpr int SIZET=42;               // sizeT = min(nV,nC)
#define WQP (w.qp)
#define QQ(i,j) w.Q[(i)+LD*(j)]       // Types.hpp:141-147: Q column-major
// R (nZ x nZ upper triangular) and T (nAC x nAC reverse triangular) are the only parts of the reference's 42x42 R and T
// arrays that are ever read, and nZ + nAC = nFR <= 42, so both live PACKED in one 903-entry buffer: T's row i (its i+1
// entries TT(i,sizeT-1-i) .. TT(i,sizeT-1)) grows from the front, R's column j (RR(0,j) .. RR(j,j)) from the back.
constexpr int TRI42=42*43/2;
WBC_HD int tri(int i){ return (i*(i+1))>>1; }
#define RR(i,j) w.RT[TRI42-1-(tri(j)+(i))]
#define TT(i,j) w.RT[tri(i)+(SIZET-1-(j))]
// Arithmetic deviations from qpOASES, all at the level of one rounding: triangular solves and DPOTRF multiply by a
// precomputed reciprocal of the diagonal where the reference divides (a divide is ~10 multiplies of latency on the critical path of every
// substitution step), the Givens chains of addConstraint / addBound come from prefix sums of squares instead of the serial hypot recurrence,
// and nvcc contracts a*b+c into FMAs.  -DWBC_EXACT (libwbc_b200_exact.so, built with -fmad=false) removes all three: that build walks the
// SAME iteration path as the reference library on every tick, including the zero-length steps whose ramping / drift-correction branch is
// decided by rounding noise (tests/test_gpu_pinned_parity.py requires 100 % equal nWSR and working sets from it).
#if defined(WBC_EXACT)
  #define WBC_EXACT_DIV
  #define WBC_EXACT_GIVENS
  #define WBC_EXACT_BACKSUB
#endif
#if defined(WBC_EXACT_DIV)
  #define WBC_RDIAG(d) (d)
  #define WBC_BYDIAG(s,r) ((s)/(r))
#else
  #define WBC_RDIAG(d) (1.0/(d))
  #define WBC_BYDIAG(s,r) ((s)*(r))
#endif
#if defined(WBC_EXACT)
  #define WBC_FNMA(l,a,s) ((s)-(l)*(a))
#else
  #define WBC_FNMA(l,a,s) fma(-(l),(a),(s))
#endif
// Exact build only: the reference's full linear-independence test (a whole step-direction solve) is run on EVERY add, next to the |Z'a|
// shortcut of the product build, and the two verdicts are compared: [0] tests, [1] tests inside the shortcut's undecided band, [2] disagreements.
#if defined(WBC_EXACT) && defined(__CUDACC__)
  __device__ unsigned long long g_wbc_li[4];
  #define WBC_LI_COUNT(k) { if(lane==0) atomicAdd(&g_wbc_li[k],1ull); }
#elif defined(WBC_EXACT)
  static unsigned long long g_wbc_li[4];
  #define WBC_LI_COUNT(k) { g_wbc_li[k]++; }
#endif
// measurement code (clock64 phase timers, algorithmic-flop counter) exists only in the profiling build (-DWBC_PROFILE -> libwbc_b200_prof.so)
#if defined(WBC_PROFILE) && defined(__CUDACC__)
  #define WBC_TIC(t) long long t=clock64()
  #define WBC_TOC(t,k) { long long n_=clock64(); c.prof[k]+=n_-t; t=n_; }
#else
  #define WBC_TIC(t) do{}while(0)
  #define WBC_TOC(t,k) do{}while(0)
#endif
#if defined(WBC_PROFILE)
  #define WBC_FLOPS(expr) { c.flops+=(expr); }
  #define WBC_COUNT_ITER() { c.prof[7]++; }
#else
  #define WBC_FLOPS(expr) do{}while(0)
  #define WBC_COUNT_ITER() do{}while(0)
#endif

// persistent solver state of one instance between ticks (HBM) -- what SQProblem keeps across hotstarts
struct QPState {
    real x[42], y[112], lb[42], ub[42], Ax_l[70], Ax_u[70];
    signed char bstat[42], cstat[70];   // working set; list ORDER is re-derived afresh by setupNewAuxiliaryQP
    int rampOffset, status, pad_;
};

// per-instance shared-memory workspace (32.4 KB in fp64 with the lane-owned vectors in tensor memory -> 7 instances per SM; 19.0 KB in fp32 -> 11).  The members up to and including `arow` are all the warm-start
// TQ rebuild touches (rebuild_TQ): the kernel that runs only that phase allocates just this prefix (QPWORK_TQ_BYTES, 9 instances per SM).
// ---------------------------------------------------------------- lane-owned vectors
// g, lb, ub, lbA, ubA, Ax, Ax - lbA, ubA - Ax, the data shift of lbA / ubA and the dual step of the fixed variables are only ever touched element by element, element i by
// lane i % 32 (loops over a list are written as loops over the element number with the status as a guard).  The device build keeps them in TENSOR MEMORY instead of
// shared memory: 256 KB per SM that nothing else in this library uses, private to a lane.  That takes the workspace of a robot from 37.6 to 32.4 KB, i.e. from
// 6 to 7 robots per SM.  Slot s of a vector = its elements 32 s .. 32 s + 31; a slot is one (fp32) or two (fp64) TMEM columns.
//   LV_FOR(i,sl,n)      loop over the elements: warp-uniform trip count with WBC_TMEM (every lane executes the tcgen05 instructions; LV_IN(i,n) guards the rest)
//   LV_LD(v,name,i,sl)  declare v and request element i (no wait); LV_WAIT(); then LV_USE(v) for every requested value before its first use
//   LV_ST(name,i,sl,v)  store (no wait); LV_DONE() at the end of the loop body
// The fp32 build keeps them in shared memory: its 11 robots per SM already saturate the instruction supply (measured: 12 robots with tensor memory = same
// speed, 13 = 10 % slower; profiles/r02_gang_variants.txt).  -DWBC_NO_TMEM: development variant of the fp64 build with the vectors in shared memory (6 robots per SM).
#if defined(__CUDACC__) && !defined(WBC_NO_TMEM) && !defined(WBC_FP32)
  #define WBC_LV_TMEM 1
#else
  #define WBC_LV_TMEM 0
#endif
// What the loops over the lane-owned vectors need from the solver context, copied out once per loop (LV_FOR declares it): the context lives in local memory
// (its address is passed to non-inlined functions) and every tcgen05.wait / st is a compiler memory barrier, so `c.tm`, `c.farbound`, ... themselves would
// be re-read (LDL) ahead of every access.
#define LV_HOIST() const LVHoist lv_h_{c.tm,c.farRampOffset,c.farbound,c.qg}
enum { LV_g=0, LV_lb=2, LV_ub=4, LV_dyFX=6, LV_lbA=8, LV_ubA=11, LV_Ax=14, LV_Ax_l=17, LV_Ax_u=20, LV_dlbA=23, LV_dubA=26, LV_SLOTS=29 };
#if WBC_LV_TMEM
  #define LV_FOR(i,sl,n) if(LV_HOIST(); true) _Pragma("unroll") for(int sl=0,i=lane;sl<((n)+31)/32;sl++,i+=32)
  #define LV_IN(i,n) ((i)<(n))
  #define LV_LD(v,name,i,sl) real v; wbc_tm_ld(lv_h_.tm+WBC_TM_COLS_PER_REAL*(unsigned)(LV_##name+(sl)),v)
  #define LV_WAIT() wbc_tm_wait_ld()
  #define LV_USE(v) wbc_tm_use(v)
  #define LV_ST(name,i,sl,v) wbc_tm_st(lv_h_.tm+WBC_TM_COLS_PER_REAL*(unsigned)(LV_##name+(sl)),(v))
  #define LV_DONE() wbc_tm_wait_st()
#else
  #define LV_FOR(i,sl,n) if(LV_HOIST(); true) LANE_FOR(i,n)
  #define LV_IN(i,n) true
  #define LV_LD(v,name,i,sl) real v=w.name[i]
  #define LV_WAIT() do{}while(0)
  #define LV_USE(v) do{}while(0)
  #define LV_ST(name,i,sl,v) w.name[i]=(v)
  #define LV_DONE() do{}while(0)
#endif
struct alignas(16) QPWork {
    real Q[42*LD];
    real RT[TRI42+1];                         // packed R and T (see RR / TT)
    signed char FR[42], FX[42], AC[70], IAC[70], bstat[42], cstat[70], posAC[70], pad_[10];     // 416 B: Q, RT, the scratch block and qp all start on 16-byte boundaries (bulk copies)
    // phase-local scratch, overlaid: Givens / row vectors + {ratio tests | step-direction temporaries} + LI-exchange vectors,
    // all of it reused as the 12 x nZ block (HZ)_force by the projected Cholesky (which runs when none of them is live)
    union {
        struct {
            real tmpv[42], arow[42];
            // three groups that are never live together: ratio tests (performStep, LI exchange) | step-direction temporaries | Givens vectors of an add
            // (computed only after the LI test / exchange of the same add has finished with the other two)
            union { struct { real num[70], den[70], dAx[70]; }; struct { real tA[42], tB[42], xTMP[42], yTMP[42], vm[42]; }; struct { real wv[42], cs_c[42], cs_s[42]; }; };
            real xFRy[42], xFRz[42], Zdz[42], ym[70];
        };
        real HZf[12*HZLD];
    };
    QPHotR qp;                                  // the matrix blocks of the compact QP (H, A); lbA / ubA / bias_j are read from HBM through QPCtx::qg
    real x[42], y[112];
#if !WBC_LV_TMEM
    // the element-wise ("lane-owned") vectors: every access is element i by the lane i % 32 -- see LV_* above; with WBC_TMEM they live in tensor memory
    real g[42], lb[42], ub[42], lbA[70], ubA[70], Ax[70], Ax_l[70], Ax_u[70];
    real dlbA[70], dubA[70];                                     // data shift of the constraint bounds (the shifts of g, lb, ub are re-evaluated where used)
    real dyFX[42];                                               // dual step of the fixed variables, by variable number
#endif
    real dx[42], dyAC[42];                                     // step (dx by variable number; dyAC by AC position)
    real rdR[42], rdT[42];                                     // reciprocal diagonals of R and T
};
constexpr size_t QPWORK_TQ_BYTES=(42*LD+TRI42+1)*sizeof(real)+416+5*42*sizeof(real);
static_assert(offsetof(QPWork,tmpv)+5*42*sizeof(real)==QPWORK_TQ_BYTES,"layout of the TQ-rebuild prefix");
static_assert(sizeof(QPWork)%16==0&&(sizeof(real)==4||offsetof(QPWork,qp)%16==0)&&offsetof(QPWork,RT)==42*LD*sizeof(real),"16-byte layout the bulk copies of the solve kernel rely on");

// what the warm-start TQ rebuild hands to the solve kernel (HBM): Q = [Z Y], the nAC rows of T, the index lists
struct alignas(16) QPPre {
    real Q[42*LD];
    real T[TRI42+1];                          // only the first tri(nAC) entries are written; Q and T have the layout of QPWork's Q and RT (one bulk copy brings both)
    signed char AC[70], IAC[70], cstat[70], posAC[70];
    int nAC, nIAC;
    double flops;
};
static_assert(sizeof(QPPre)%16==0&&offsetof(QPPre,T)==42*LD*sizeof(real),"QPPre: Q and T contiguous like QPWork's Q and RT, records 16-byte aligned");

struct LVHoist { unsigned tm; int farRampOffset; real farbound; const QPData* qg; };
// scalar solver state, replicated in every lane
// The context is reached by reference from non-inlined functions, so it lives in local memory and every field read is an LDL; the lane number at least
// comes from the thread index instead (every kernel that runs the solver has lane = threadIdx.x % 32).
#if defined(__CUDACC__)
  #define WBC_LANE(c) ((int)(threadIdx.x&31u))
#else
  #define WBC_LANE(c) ((c).lane)
#endif
struct QPCtx {
    int lane;
    int nFR,nFX,nAC,nIAC;
    int status, infeasible, unbounded, rampOffset;
    int frIdentity;      // 1 while the free list is 0,1,..,41 in this order (every warm start until a bound becomes active): products with Q skip the list indirection
    int farRampOffset;   // rampOffset at the last updateFarBounds
    real farbound;     // far bound of the current homotopy (QProblem::updateFarBounds)
    real tau;
    const QPData* qg;    // this robot's compact QP in HBM (fp64): new lbA / ubA targets, bias_j
    unsigned tm;                // WBC_TMEM: tensor-memory address (lane quarter of this warp, first column of its robot) of the lane-owned vectors
    unsigned tqBar, tqParity;   // != 0: a bulk copy of the rebuilt Q / T into w.Q, w.RT is in flight on this mbarrier (shared-space address) with this phase parity
    // parking (wbc_solve_cold_kernel: cold batches run gang by gang; when the first member of a gang is done, the others stop at the top of their next iteration, their
    // whole state goes to HBM and a second launch finishes them with per-warp fetching -- no warp idles while the slowest member of its gang finishes)
    int parkOK;                                          // 1: this launch may park the solve (phase 1); 0 in the launch that finishes the parked ones
    int resuming;                                        // 1: the state was restored from a parked solve; qp_init / hotstartFar / solveQP skip their prologues once
    int park_iter, park_nWSRmax; real park_farbound;     // loop state of solveQP / hotstartFar at the suspension
#if defined(WBC_PROFILE)
    long long prof[8];   // clock64 ticks: 0 data shift, 1 step direction, 2 perform step, 3 change active set, 4 projected Cholesky, 5 ramp/drift, 6 whole solve, 7 iterations
    double flops;        // ALGORITHMIC flops of the qpOASES method on dense data (SURVEY.md section 8d formulas), not what this kernel executes
#endif
};

// a parked solve: the whole workspace, the lane-owned vectors (tensor-memory build) and the context
struct alignas(16) QPPark {
    QPWork w;
    real lv[LV_SLOTS*32];
    QPCtx c;
};
static_assert(offsetof(QPPark,w)==0&&sizeof(QPWork)%16==0,"QPPark: the workspace is restored by one bulk copy");

// ---------------------------------------------------------------- gang mode (WBC_GANG = warps per CTA, one robot per warp)
// The solver is instruction-SUPPLY bound when its warps run independently: 193 KB of code, ~90 KB touched per working-set change, a 32 KB
// L1.5 instruction cache per SM and a GPC-level cache that delivers one line per two cycles to ~18 SMs (ncu, 5 independent warps per SM:
// sm__icc hit rate 62 %, gcc instruction requests 85 % of peak, 2.5 of 7.6 stall cycles per instruction are `no_instruction`).
// In gang mode the WBC_GANG warps of a CTA meet at ONE barrier at the top of every working-set iteration, so they walk the phases of the
// iteration roughly together and share the instruction-cache fills (icc hit rate 91 %, gcc requests 38 % of peak, no_instruction 0.26).
// More barrier sites per iteration were measured and do not pay (profiles/r02_gang_variants.txt).  The barrier also reduces a predicate
// (every warp done?) which the drain loops of the kernel use to leave together.
#if defined(WBC_GANG) && defined(__CUDACC__)
  WBC_DEV bool gang_bar(int done){ return __syncthreads_and(done)!=0; }
  // the same barrier with a count instead of a predicate (how many threads of the gang are done); a launch uses ONE of the two forms at every site
  WBC_DEV int gang_count(int done){ return __syncthreads_count(done); }
  #define WBC_GANG_SYNC() gang_bar(0)
#else
  #define WBC_GANG_SYNC() do{}while(0)
#endif

// ---------------------------------------------------------------- structured A and H
// element decoders in the type of the data they read (fp64 for the host-side dense-QP getter, `real` for the solver's shared-memory copy)
template<class QD> WBC_HD typename QD::value_type Ael(const QD& q,int i,int j){
    typedef typename QD::value_type T;
    if(i<6){ if(j<6) return q.Mbb[6*i+j]; if(j<18||j>=30) return T(0.0); int r=j-18; return -(q.cmask[r/3]*q.Jc[18*r+i]); }
    if(i<18){ int r=i-6; return (j<18)? q.cmask[r/3]*q.Jc[18*r+j] : T(0.0); }
    if(i<34){ int r=i-18,l=r>>2,k=r&3,c=j-18-3*l; if(c<0||c>2) return T(0.0);
              T d = (c==2)? -q.mu : ((c==(k>>1))? ((k&1)?T(-1.0):T(1.0)) : T(0.0)); return d*q.cmask[l]; }
    if(i<46){ int r=i-34; if(j<6||j>=30) return T(0.0); if(j<18) return q.Mjj[12*r+j-6]; int f=j-18; return -(q.cmask[f/3]*q.Jc[18*f+6+r]); }
    { int r=(i<58)?i-46:i-58; if(j<18) return (T(1.0)-q.cmask[r/3])*q.Jc[18*r+j]; if(j==30+r) return (i<58)?T(-1.0):T(1.0); return T(0.0); }
}
template<class QD> WBC_HD typename QD::value_type Hel(const QD& q,int i,int j){
    typedef typename QD::value_type T;
    if(i>=18&&i<30&&j>=18&&j<30) return q.Hff[12*(i-18)+(j-18)];
    if(i!=j) return T(0.0); return (i>=30)?q.rho:T(1.0);
}
template<class QD> WBC_FN real Ael_dev(const QD& q,int i,int j){ return (real)Ael(q,i,j); }       // ONE device copy of the element decoder
// Structured products with A (whole_body_controller.cpp:492-497): no per-element decoding, only the non-zero blocks.
// All of them ACCUMULATE ONTO THE GIVEN VALUE term by term, y <- y + sgn*a_ij*v_j in ascending j, which is what the reference's dense
// loops do (DenseMatrix::times / transTimes with index lists, Matrices.cpp:332-590, and the DGEMM replacement, BLASReplacement.cpp:36-92:
// `y[.] += val*x` / `y[.] -= val*x` onto the existing y; zero entries of the dense row add exactly 0).  sgn is +1.0 or -1.0 exactly,
// so (sgn*a)*v == +-(a*v) bit for bit.
// Arow_acc: y + sgn*(A v)_i for a 42-vector v (callers mask v to the free / fixed columns beforehand: the reference visits the selected columns only).
WBC_FN real Arow_acc(const QPHotR& q,int i,const real* v,real y,real sgn){
    if(i<6){ for(int j=0;j<6;j++) y+=(sgn*q.Mbb[6*i+j])*v[j]; for(int f=0;f<12;f++) y-=(sgn*(q.cmask[f/3]*q.Jc[18*f+i]))*v[18+f]; return y; }
    if(i<18){ const int r=i-6; const real c=q.cmask[r/3]; if(c==real(0.0)) return y; for(int j=0;j<18;j++) y+=(sgn*q.Jc[18*r+j])*v[j]; return y; }
    if(i<34){ const int r=i-18,l=r>>2,k=r&3; const real c=q.cmask[l]; if(c==real(0.0)) return y; const real* f=v+18+3*l;
              const real d=(k&1)?-sgn:sgn;
              if(k<2){ y+=d*f[0]; } else { y+=d*f[1]; }
              return y-(sgn*q.mu)*f[2]; }
    if(i<46){ const int r=i-34; for(int j=0;j<12;j++) y+=(sgn*q.Mjj[12*r+j])*v[6+j]; for(int f=0;f<12;f++) y-=(sgn*(q.cmask[f/3]*q.Jc[18*f+6+r]))*v[18+f]; return y; }
    { const int r=(i<58)?i-46:i-58; const real c=real(1.0)-q.cmask[r/3]; if(c!=real(0.0)) for(int j=0;j<18;j++) y+=(sgn*q.Jc[18*r+j])*v[j]; return (i<58)? y-sgn*v[30+r] : y+sgn*v[30+r]; }
}
// Acol_acc: y + sgn*(A' yv)_j for a 70-vector yv by constraint number (0 for inactive constraints), rows visited in ascending order
WBC_FN real Acol_acc(const QPHotR& q,int j,const real* yv,real y,real sgn){
    if(j<18){
        if(j<6) for(int i=0;i<6;i++) y+=(sgn*q.Mbb[6*i+j])*yv[i];
        for(int r=0;r<12;r++) y+=(sgn*(q.cmask[r/3]*q.Jc[18*r+j]))*yv[6+r];
        if(j>=6) for(int r=0;r<12;r++) y+=(sgn*q.Mjj[12*r+(j-6)])*yv[34+r];
        for(int r=0;r<12;r++) y+=(sgn*((real(1.0)-q.cmask[r/3])*q.Jc[18*r+j]))*yv[46+r];
        for(int r=0;r<12;r++) y+=(sgn*((real(1.0)-q.cmask[r/3])*q.Jc[18*r+j]))*yv[58+r];
        return y;
    }
    if(j<30){ const int f=j-18,l=f/3,c=f-3*l; const real cf=q.cmask[l];
        for(int i=0;i<6;i++) y-=(sgn*(cf*q.Jc[18*f+i]))*yv[i];
        const real* yf=yv+18+4*l;
        if(c==2){ const real m=sgn*(cf*q.mu); y-=m*yf[0]; y-=m*yf[1]; y-=m*yf[2]; y-=m*yf[3]; }
        else { const real d=sgn*cf; y+=d*yf[2*c]; y-=d*yf[2*c+1]; }
        for(int r=0;r<12;r++) y-=(sgn*(cf*q.Jc[18*f+6+r]))*yv[34+r];
        return y; }
    y-=sgn*yv[46+(j-30)];
    return y+sgn*yv[58+(j-30)];
}
// (H v)_i over the free (sel 0), fixed (sel 1) or all (sel 2) variables.  H = blkdiag(I6, I12, Hff, rho I12): only the 12 force rows
// have off-diagonal terms; `fm` = bit mask of the force variables selected by `sel` (force_mask), terms are added in ascending
// variable number like the reference's index-list products (Matrices.cpp:384-397).  Hff is bitwise symmetric: read by column.
WBC_DEV unsigned force_mask(const QPWork& w,const QPCtx& c,int sel){
#if WBC_NL==1
    (void)c; unsigned m=0u; for(int u=0;u<12;u++) if(sel==2||((w.bstat[18+u]!=0)==(sel==1))) m|=1u<<u; return m;
#else
    return wbc_ballot(WBC_LANE(c)<12&&(sel==2||((w.bstat[18+(WBC_LANE(c)<12?WBC_LANE(c):0)]!=0)==(sel==1))));
#endif
}
WBC_FN real Hrow_acc(const QPWork& w,int i,const real* v,int sel,unsigned fm,real y,real sgn){
    if(i>=18&&i<30){ const real* h=&WQP.Hff[i-18];
        #pragma unroll
        for(int u=0;u<12;u++) if((fm>>u)&1u) y+=(sgn*h[12*u])*v[18+u];
        return y; }
    if(sel!=2 && ((w.bstat[i]!=0)!=(sel==1))) return y;
    return y+(sgn*((i>=30)?WQP.rho:real(1.0)))*v[i];
}
// masked copies: vm = v restricted to free (sel 0) / fixed (sel 1) variables; ym = yAC scattered to constraint numbers
WBC_FN void mask_vars(QPWork& w,const QPCtx& c,const real* v,int sel){ const int lane=WBC_LANE(c); WBC_SYNC(); LANE_FOR(j,42) w.vm[j]=((w.bstat[j]!=0)==(sel==1))?v[j]:real(0.0); WBC_SYNC(); }
WBC_FN void scatter_yAC(QPWork& w,const QPCtx& c,const real* yAC){ const int lane=WBC_LANE(c); WBC_SYNC(); LANE_FOR(i,70) w.ym[i]=(w.cstat[i]!=0)?yAC[(int)w.posAC[i]]:real(0.0); WBC_SYNC(); }

// ---------------------------------------------------------------- products with Q = [Z Y] (one copy each: instruction cache)
// out[i] = sgn * sum_j Q(FR[j], col0+i) t[FR[j]],  i < ncols   (t by variable number; lane per column; FR-list order like the reference).
// A lane owns columns i and i+32 and walks both in the same loop (two independent chains, t loaded once).  While the free list is
// the identity (c.frIdentity) the list indirection is skipped -- same terms, same order.
WBC_FN void Qt_times(const QPWork& w,const QPCtx& c,int col0,int ncols,const real* t,real sgn,real* out){
    const int lane=WBC_LANE(c), nFR=c.nFR;
#if WBC_NL==1
    for(int i=0;i<ncols;i++){ const real* q=&w.Q[LD*(col0+i)]; real s=real(0.0); for(int j=0;j<nFR;j++){ const int jj=(int)w.FR[j]; s+=q[jj]*t[jj]; } out[i]=sgn*s; }
#else
    const bool in0=lane<ncols, in1=lane+32<ncols;
    const real* q0=&w.Q[LD*(col0+(in0?lane:0))]; const real* q1=&w.Q[LD*(col0+(in1?lane+32:0))];
    real s0=real(0.0),s1=real(0.0);
    if(c.frIdentity){
        if(ncols>32){
            #pragma unroll 6
            for(int j=0;j<42;j++){ const real tj=t[j]; s0+=q0[j]*tj; s1+=q1[j]*tj; }
        } else {
            #pragma unroll 6
            for(int j=0;j<42;j++) s0+=q0[j]*t[j];
        }
    } else if(ncols>32){
        #pragma unroll 4
        for(int j=0;j<nFR;j++){ const int jj=(int)w.FR[j]; const real tj=t[jj]; s0+=q0[jj]*tj; s1+=q1[jj]*tj; }
    } else {
        #pragma unroll 4
        for(int j=0;j<nFR;j++){ const int jj=(int)w.FR[j]; s0+=q0[jj]*t[jj]; }
    }
    if(in0) out[lane]=sgn*s0;
    if(in1) out[lane+32]=sgn*s1;
#endif
    WBC_SYNC();
}
// v = sum_j Q(FR[i], col0+j) x[j] for every free variable FR[i]:  acc[FR[i]] += v,  and keep[FR[i]] = v if keep != null   (lane per row,
// rows i and i+32 of the list in one loop)
WBC_FN void Q_times(const QPWork& w,const QPCtx& c,int col0,int ncols,const real* x,real* acc,real* keep){
    const int lane=WBC_LANE(c), nFR=c.nFR;
#if WBC_NL==1
    for(int i=0;i<nFR;i++){ const int ii=(int)w.FR[i]; const real* q=&w.Q[ii+LD*col0]; real s=real(0.0); for(int j=0;j<ncols;j++) s+=q[LD*j]*x[j]; if(keep) keep[ii]=s; acc[ii]+=s; }
#else
    const bool in0=lane<nFR, in1=lane+32<nFR;
    const int i0=in0?(int)w.FR[lane]:0, i1=in1?(int)w.FR[lane+32]:0;
    const real* q0=&w.Q[i0+LD*col0]; const real* q1=&w.Q[i1+LD*col0];
    real s0=real(0.0),s1=real(0.0);
    if(nFR>32){
        #pragma unroll 4
        for(int j=0;j<ncols;j++){ const real xj=x[j]; s0+=q0[LD*j]*xj; s1+=q1[LD*j]*xj; }
    } else {
        #pragma unroll 4
        for(int j=0;j<ncols;j++) s0+=q0[LD*j]*x[j];
    }
    if(in0){ if(keep) keep[i0]=s0; acc[i0]+=s0; }
    if(in1){ if(keep) keep[i1]=s1; acc[i1]+=s1; }
#endif
    WBC_SYNC();
}

// ---------------------------------------------------------------- index lists (Indexlist.cpp: insertion-ordered number[])
// removal keeps the order of the remaining entries; every lane holds the same counters, one lane edits the list
WBC_DEV void list_remove(signed char* L,int& n,int number,int lane){
#if WBC_NL==1
    { int k=0; while(k<n&&L[k]!=number) k++; for(;k<n-1;k++) L[k]=L[k+1]; }
#else
    int k=n;                                                    // position of `number`: ballot over chunks of 32
    for(int base=0;base<n;base+=32){ const int i=base+lane; const unsigned m=wbc_ballot(i<n&&(int)L[i]==number); if(m){ k=base+wbc_ffs(m); break; } }
    signed char v[3]; bool mv[3];
    #pragma unroll
    for(int u=0;u<3;u++){ const int i=lane+32*u; mv[u]=(i>=k&&i<n-1); v[u]=mv[u]?L[i+1]:(signed char)0; }
    WBC_SYNC();
    #pragma unroll
    for(int u=0;u<3;u++) if(mv[u]) L[lane+32*u]=v[u];
#endif
    n--; WBC_SYNC();
}
WBC_DEV void list_append(signed char* L,int& n,int number,int lane){ if(lane==0) L[n]=(signed char)number; n++; WBC_SYNC(); }
WBC_DEV void rebuild_posAC(QPWork& w,const QPCtx& c){ const int lane=WBC_LANE(c); LANE_FOR(k,c.nAC) w.posAC[(int)w.AC[k]]=(signed char)k; WBC_SYNC(); }

// Bounds::moveFreeToFixed / moveFixedToFree, Constraints::moveInactiveToActive / moveActiveToInactive (Bounds.cpp, Constraints.cpp)
WBC_DEV void moveFreeToFixed(QPWork& w,QPCtx& c,int number,int st){ c.frIdentity=0; WBC_SYNC(); list_remove(w.FR,c.nFR,number,WBC_LANE(c)); list_append(w.FX,c.nFX,number,WBC_LANE(c)); if(WBC_LANE(c)==0) w.bstat[number]=(signed char)st; WBC_SYNC(); }
WBC_DEV void moveFixedToFree(QPWork& w,QPCtx& c,int number){ c.frIdentity=0; WBC_SYNC(); list_remove(w.FX,c.nFX,number,WBC_LANE(c)); list_append(w.FR,c.nFR,number,WBC_LANE(c)); if(WBC_LANE(c)==0) w.bstat[number]=0; WBC_SYNC(); }
WBC_DEV void moveInactiveToActive(QPWork& w,QPCtx& c,int number,int st){ WBC_SYNC(); list_remove(w.IAC,c.nIAC,number,WBC_LANE(c)); if(WBC_LANE(c)==0){ w.cstat[number]=(signed char)st; w.posAC[number]=(signed char)c.nAC; } list_append(w.AC,c.nAC,number,WBC_LANE(c)); }
WBC_DEV void moveActiveToInactive(QPWork& w,QPCtx& c,int number){ WBC_SYNC(); list_remove(w.AC,c.nAC,number,WBC_LANE(c)); list_append(w.IAC,c.nIAC,number,WBC_LANE(c)); if(WBC_LANE(c)==0) w.cstat[number]=0; WBC_SYNC(); rebuild_posAC(w,c); }
// Indexlist::swapNumbers on the free list (Bounds::swapFree)
WBC_DEV void swapFree(QPWork& w,QPCtx& c,int n1,int n2){ c.frIdentity=0; WBC_SYNC(); if(WBC_LANE(c)==0){ int a=0,b=0; while(w.FR[a]!=n1) a++; while(w.FR[b]!=n2) b++; signed char t=w.FR[a]; w.FR[a]=w.FR[b]; w.FR[b]=t; } WBC_SYNC(); }

// ---------------------------------------------------------------- triangular solves
// Forward substitution L u = b,  u_k = (b[k*bs] - sum_{j<k} L(k,j) u_j) * rd[k*rs],  a[k*as] = u_k, for the four triangular
// systems of the method (QProblem::backsolveT :4518-4565, QProblemB::backsolveR :1635-1688) on the packed R/T buffer:
//   kind 0: T    L(i,j) = TT(i,sizeT-1-j)          kind 1: T'   L(i,j) = TT(n-1-j,sizeT-n+i)
//   kind 2: R'   L(i,j) = RR(j,i)                  kind 3: R    L(i,j) = RR(n-1-i,n-1-j)
// In packed storage the address of L(i,j+1) is that of L(i,j) plus d0 + j*dd with (d0,dd) = (1,0), (-(n-1),1), (-1,0), (n-1,-1),
// so ONE routine (one copy of the code: the solver is instruction-cache bound) walks all four.
// Column-oriented: lane r owns row r (and r+32), keeps its running sum in a register and subtracts the terms in
// ascending j, the order of the reference's row loops; u_k travels by one shuffle.  Diagonals enter as precomputed
// reciprocals (fp64 division is ~20x the latency of a multiply and sat on the critical path of every step).
WBC_DEV int trisolve_row0(int kind,int n,int i){      // packed index of L(i,0)
    return kind==0? tri(i) : kind==1? tri(n-1)+n-1-i : kind==2? TRI42-1-tri(i) : TRI42-1-(tri(n-1)+n-1-i);
}
WBC_FN void trisolve(int lane,int kind,int n,const real* RT,const real* b,int bs,real* a,int as,const real* rd,int rs){
    const int d0= kind==0? 1 : kind==1? -(n-1) : kind==2? -1 : n-1, dd= kind==1? 1 : kind==3? -1 : 0;
#if WBC_NL==1
    (void)lane; real u[42];
    for(int k=0;k<n;k++){ real sum=b[k*bs]; int off=trisolve_row0(kind,n,k), d=d0;
        for(int j=0;j<k;j++){ sum-=RT[off]*u[j]; off+=d; d+=dd; }
        u[k]=WBC_BYDIAG(sum,rd[k*rs]); }
    for(int k=0;k<n;k++) a[k*as]=u[k];
#else
    const bool in0=lane<n, in1=lane+32<n;
    const unsigned L=wbc_saddr(RT);                           // 32-bit shared address of the packed R/T buffer (read-only here)
    int o0=in0?trisolve_row0(kind,n,lane):0, d=d0;
    real s0=in0?b[lane*bs]:real(0.0), r0=in0?rd[lane*rs]:real(1.0), l0=(in0&&lane>0)?wbc_lds(L+(unsigned)sizeof(real)*o0):real(0.0);
    WBC_SYNC();                                              // a may alias b
    // l0 is loaded as 0 for lanes <= k, so the update needs no predicate and a finished lane keeps its final sum: u_k = s0*r0 is
    // formed for the broadcast at step k and once more, from the same operands, for the store after the loop.
    if(n<=32){
        #pragma unroll 2
        for(int k=0;k<n;k++){
            const real ak=__shfl_sync(0xffffffffu,WBC_BYDIAG(s0,r0),k);
            o0+=d; d+=dd;
            const real nl=(lane>k+1&&in0)?wbc_lds(L+(unsigned)sizeof(real)*o0):real(0.0);   // next column's entry, independent of the chain
            s0=WBC_FNMA(l0,ak,s0);
            l0=nl;
        }
        if(in0) a[lane*as]=WBC_BYDIAG(s0,r0);
    } else {
        int o1=in1?trisolve_row0(kind,n,lane+32):0;
        real s1=in1?b[(lane+32)*bs]:real(0.0), r1=in1?rd[(lane+32)*rs]:real(1.0), l1=in1?wbc_lds(L+(unsigned)sizeof(real)*o1):real(0.0);
        #pragma unroll 2
        for(int k=0;k<n;k++){
            const real mine=(k<32)?WBC_BYDIAG(s0,r0):WBC_BYDIAG(s1,r1);
            const real ak=__shfl_sync(0xffffffffu,mine,k&31);
            o0+=d; o1+=d; d+=dd;
            const real nl0=(lane>k+1&&in0)?wbc_lds(L+(unsigned)sizeof(real)*o0):real(0.0), nl1=(lane+32>k+1&&in1)?wbc_lds(L+(unsigned)sizeof(real)*o1):real(0.0);
            s0=WBC_FNMA(l0,ak,s0); s1=WBC_FNMA(l1,ak,s1);
            l0=nl0; l1=nl1;
        }
        a[lane*as]=WBC_BYDIAG(s0,r0);
        if(in1) a[(lane+32)*as]=WBC_BYDIAG(s1,r1);
    }
#endif
    WBC_SYNC();
}
// reciprocals of T's diagonal TT(k, sizeT-1-k); false if a pivot fails the |.| > EPS test of QProblem::backsolveT
WBC_FN bool prepare_T(QPWork& w,const QPCtx& c){
    const int lane=WBC_LANE(c); bool bad=false; WBC_SYNC();
    LANE_FOR(k,c.nAC){ const real d=TT(k,SIZET-1-k); if(!(fabs(d)>QEPS)) bad=true; w.rdT[k]=WBC_RDIAG(d); }
    WBC_SYNC();
    return wbc_ballot(bad)==0u;
}
// QProblem::backsolveT (QProblem.cpp:4518-4565).  b, a indexed by AC position.  prepare_T must have run since T last changed.
// unknown u_k = a[nT-1-k]; row i, column j<i:  TT(i,sizeT-1-j)  resp.  TT(nT-1-j,tcol+i) for the transposed system
WBC_DEV void backsolveT(QPWork& w,const QPCtx& c,const real* b,bool transposed,real* a){
    const int nT=c.nAC;
    if(nT<=0) return;
    if(!transposed) trisolve(WBC_LANE(c),0,nT,w.RT,b,1,a+nT-1,-1,w.rdT,1);
    else            trisolve(WBC_LANE(c),1,nT,w.RT,b,1,a+nT-1,-1,w.rdT+nT-1,-1);
}
// QProblemB::backsolveR (QProblemB.cpp:1635-1688) on the nR x nR upper-triangular R; a may alias b.  w.rdR = 1/diag(R).
WBC_DEV void backsolveR(QPWork& w,const QPCtx& c,const real* b,bool transposed,real* a,int nR){
    if(nR<=0) return;
    if(transposed){ trisolve(WBC_LANE(c),2,nR,w.RT,b,1,a,1,w.rdR,1); return; }
#if defined(WBC_EXACT_BACKSUB)
    // The reference's back substitution subtracts the terms of row i in ASCENDING column order j = i+1 .. nR-1, i.e. starting with the unknown
    // it has only just computed: a serial chain of nR^2/2 operations that the streaming form below (terms in the order the unknowns become
    // available: j = nR-1 down to i+1) re-orders.  Exact build: the serial form, every lane redundantly, lane 0 stores.
    { const int lane=WBC_LANE(c); WBC_SYNC();
      for(int i=nR-1;i>=0;--i){ real sum=b[i]; for(int j=i+1;j<nR;++j) sum-=RR(i,j)*a[j]; const real v=WBC_BYDIAG(sum,w.rdR[i]); WBC_SYNC(); if(lane==0) a[i]=v; WBC_SYNC(); } }
#else
    trisolve(WBC_LANE(c),3,nR,w.RT,b+nR-1,-1,a+nR-1,-1,w.rdR+nR-1,-1);
#endif
}

// ---------------------------------------------------------------- factorisations
// QProblem::setupTQfactorisation (QProblem.cpp:2169-2193)
WBC_DEV void setupTQfactorisation(QPWork& w,const QPCtx& c){
    const int lane=WBC_LANE(c); WBC_SYNC();
    LANE_FOR(e,42*LD) w.Q[e]=real(0.0);                     // packed T needs no clearing: every entry of a row is written when the row is created
    WBC_SYNC();
    LANE_FOR(i,c.nFR) QQ((int)w.FR[i],i)=real(1.0);
    WBC_SYNC();
}
// DPOTRF "U" of LAPACKReplacement.cpp:39-77 on the leading n x n block of R.  Row i: the lane that owns column j forms R(i,j); its
// inner product runs k = i-1 .. 0 like the reference; the diagonal's sum travels by one shuffle from the lane that owns column i.
// returns false if not SPD.
WBC_DEV bool potrf_R(QPWork& w,const QPCtx& c,int n){
    const int lane=WBC_LANE(c);
    WBC_SYNC();
#if WBC_NL==1
    for(int i=0;i<n;i++){
        real sum=RR(i,i); for(int k=i-1;k>=0;k--) sum-=RR(k,i)*RR(k,i);
        if(!(sum>real(0.0))) return false;
        const real d=sqrt(sum), rd=WBC_RDIAG(d); RR(i,i)=d; w.rdR[i]=rd;
        for(int j=i+1;j<n;j++){ real s2=RR(i,j); for(int k=i-1;k>=0;k--) s2-=RR(k,i)*RR(k,j); RR(i,j)=WBC_BYDIAG(s2,rd); }
    }
#else
    // lane t owns columns t and t+32 for the whole factorisation (fixed pointers, no per-row index arithmetic); packed column j lives
    // at RT[TRI42-1-tri(j) - row].  At row i the lanes left of the diagonal run the same loop on in-bounds data and discard the result.
    const bool own0=lane<n, own1=lane+32<n;
    real* c0=&w.RT[TRI42-1-tri(own0?lane:0)];                // &RR(0,lane)
    real* c1=&w.RT[TRI42-1-tri(own1?lane+32:0)];
    const real* ci=&w.RT[TRI42-1];                           // &RR(0,i)
    for(int i=0;i<n;i++){
        real s0=c0[-i], s1=real(0.0);
        if(n>32){
            s1=c1[-i];
            #pragma unroll 2
            for(int r=i-1;r>=0;r--){ const real rki=ci[-r]; s0-=rki*c0[-r]; s1-=rki*c1[-r]; }
        } else {
            #pragma unroll 2
            for(int r=i-1;r>=0;r--) s0-=ci[-r]*c0[-r];
        }
        const real sum=wbc_bcast((i<32)?s0:s1,i&31);
        if(!(sum>real(0.0))) return false;
        const real d=sqrt(sum), rd=WBC_RDIAG(d);
        if(lane==(i&31)){ if(i<32) c0[-i]=d; else c1[-i]=d; w.rdR[i]=rd; }
        if(own0&&lane>i) c0[-i]=WBC_BYDIAG(s0,rd);
        if(own1&&lane+32>i) c1[-i]=WBC_BYDIAG(s1,rd);
        ci-=i+1;                                                // tri(i+1)-tri(i)
        WBC_SYNC();
    }
#endif
    return true;
}
// R(kk,jj) = sum_ii Z(FR[ii],jj) (HZ)(FR[ii],kk) for the upper triangle, register-tiled: each lane accumulates 4 x 4 tiles (the
// upper tiles of the tile grid, <= 66 of them -> <= 3 per lane) over the free variables in list order -- same products and same
// summation order as the scalar loops of SymDenseMat::bilinear (Matrices.cpp:630-667).
WBC_DEV void zhz_tiles(QPWork& w,const QPCtx& c,int nFR,int nZ){
    constexpr int TS=4;
    const int lane=WBC_LANE(c), nt=(nZ+TS-1)/TS, ntiles=nt*(nt+1)/2;
    LANE_FOR(t,ntiles){
        int ti=0, rem=t; while(rem>=nt-ti){ rem-=nt-ti; ti++; } const int tj=ti+rem;
        real acc[TS][TS];
        #pragma unroll
        for(int a=0;a<TS;a++){
            #pragma unroll
            for(int b=0;b<TS;b++) acc[a][b]=real(0.0); }
        #pragma unroll 1
        for(int ii=0;ii<nFR;ii++){
            const int v=(int)w.FR[ii]; real zj[TS],wk[TS];
            #pragma unroll
            for(int b=0;b<TS;b++) zj[b]=QQ(v,TS*tj+b);
            if(v>=18&&v<30){
                #pragma unroll
                for(int a=0;a<TS;a++) wk[a]=w.HZf[(v-18)*HZLD+TS*ti+a];
            } else { const real dv=(v>=30)?WQP.rho:real(1.0);
                #pragma unroll
                for(int a=0;a<TS;a++) wk[a]=dv*QQ(v,TS*ti+a); }
            #pragma unroll
            for(int a=0;a<TS;a++){
                #pragma unroll
                for(int b=0;b<TS;b++) acc[a][b]+=zj[b]*wk[a]; }
        }
        #pragma unroll
        for(int a=0;a<TS;a++){
            #pragma unroll
            for(int b=0;b<TS;b++){ const int kk=TS*ti+a, jj=TS*tj+b; if(kk<=jj&&jj<nZ) RR(kk,jj)=acc[a][b]; } }
    }
}
// the two cases without an active constraint (first iterations of a cold start): R'R = H_FR,FR directly
WBC_FN void fill_R_with_H(QPWork& w,const QPCtx& c){
    const int lane=WBC_LANE(c), nFR=c.nFR;
    LANE_FOR(j,nFR) for(int i=0;i<=j;i++) RR(i,j)=Hel(WQP,(int)w.FR[i],(int)w.FR[j]);
}
// QProblem::computeProjectedCholesky (QProblem.cpp:2036-2129) incl. the QProblemB::computeCholesky fallback (:1366-1446)
WBC_FN int computeProjectedCholesky(QPWork& w,QPCtx& c){
    const int lane=WBC_LANE(c), nFR=c.nFR, nZ=c.nFR-c.nAC;
    WBC_FLOPS(2.0*nFR*nFR*nZ+2.0*nFR*nZ*nZ+(double)nZ*nZ*nZ/3.0);
    WBC_SYNC();
    if(nZ==0&&c.nFX+c.nAC!=0) return QP_OK;
    if(c.nAC==0){
        if(c.nFX!=0){ // make Z trivial
            LANE_FOR(i,42) for(int j=0;j<nZ;j++) QQ(i,j)=real(0.0);
            WBC_SYNC();
            LANE_FOR(j,nZ) QQ((int)w.FR[j],j)=real(1.0);
        }
        fill_R_with_H(w,c);
    } else {
        // Z'HZ (SymDenseMat::bilinear, Matrices.cpp:630-667).  H = diag except the 12x12 force block, so
        // (HZ)(v,k) = H_vv Z(v,k) for v outside the block; inside, the sum runs over the free force variables in list order
        // (w.rdT = that sub-list, built with ballots; prepare_T rewrites rdT before its next use).  Hff is bitwise symmetric, read as Hff[u][a] for unit stride in a.
        int nff=0; int* ffl=reinterpret_cast<int*>(w.rdT);      // <= 12 ints in the (currently dead) rdT buffer
        for(int base=0;base<nFR;base+=WBC_NL){ const int i=base+lane; const int v=(i<nFR)?(int)w.FR[i]:0; const bool isf=(i<nFR)&&v>=18&&v<30;
            const unsigned m=wbc_ballot(isf); if(isf) ffl[nff+wbc_rank(m,lane)]=v-18; nff+=wbc_popc(m); }
        WBC_SYNC();
        LANE_FOR(e,12*nZ){ const int k=e/12, a=e-12*k; real s=0;
            for(int i=0;i<nff;i++){ const int u=ffl[i]; s+=WQP.Hff[12*u+a]*QQ(18+u,k); }
            w.HZf[a*HZLD+k]=s; }
        WBC_SYNC();
        zhz_tiles(w,c,nFR,nZ);
    }
    if(!potrf_R(w,c,nZ)) return QP_RET_HESSIAN_NOT_SPD;
    return QP_OK;
}
// The same factor after a REMOVAL (removeBound / removeConstraint) without recomputing what cannot have changed.  A removal appends ONE column to Z
// (index n = nZ-1; the Givens sweeps of QProblem.cpp:3846-3870, 4099-4121 touch columns >= n only) and leaves the first n columns bit for bit alone; a freed
// variable joins the END of the free list with zeros in those columns, so its terms add exactly 0 to every old entry of Z'HZ.  DPOTRF "U" computes R row by
// row from the rows above, hence the leading n x n block of the new R is bitwise the old one and only column n is new:
//     A(i,n) = sum_ii Z(ii,n) (HZ)(ii,i)            (the reference's order: SymDenseMat::bilinear, free-list order)
//     R(i,n) = (A(i,n) - sum_{k=i-1..0} R(k,i) R(k,n)) / R(i,i),   R(n,n) = sqrt(A(n,n) - sum_{k=n-1..0} R(k,n)^2)      (LAPACKReplacement.cpp:39-77)
// The inner sums start with the entry computed last, so the column is one serial chain; every lane runs it redundantly on broadcast loads and stores the
// same values (a lane only ever reads back what it wrote itself: no synchronisation inside the chain).  Saves the (nZ^2/2) x nFR tile products of Z'HZ and
// nZ-1 square roots / divisions on more than half of the iterations of a cold start.  Needs nAC > 0 (with nAC == 0 the reference re-trivialises Z).
WBC_FN int extendProjectedCholesky(QPWork& w,QPCtx& c){
    const int lane=WBC_LANE(c), nFR=c.nFR, nZ=c.nFR-c.nAC, n=nZ-1;
    WBC_FLOPS(2.0*nFR*nFR*nZ+2.0*nFR*nZ*nZ+(double)nZ*nZ*nZ/3.0);      // algorithmic count of the reference, which recomputes everything
    WBC_SYNC();
    int nff=0; int* ffl=reinterpret_cast<int*>(w.rdT);
    for(int base=0;base<nFR;base+=WBC_NL){ const int i=base+lane; const int v=(i<nFR)?(int)w.FR[i]:0; const bool isf=(i<nFR)&&v>=18&&v<30;
        const unsigned m=wbc_ballot(isf); if(isf) ffl[nff+wbc_rank(m,lane)]=v-18; nff+=wbc_popc(m); }
    WBC_SYNC();
    LANE_FOR(e,12*nZ){ const int k=e/12, a=e-12*k; real s=0;
        for(int i=0;i<nff;i++){ const int u=ffl[i]; s+=WQP.Hff[12*u+a]*QQ(18+u,k); }
        w.HZf[a*HZLD+k]=s; }
    WBC_SYNC();
    real* col=w.dx;                                           // the step direction is dead between performStep and the next determineStepDirection
    LANE_FOR(i,nZ){ real s=real(0.0);
        for(int ii=0;ii<nFR;ii++){ const int v=(int)w.FR[ii]; const real z=QQ(v,n);
            const real hk=(v>=18&&v<30)? w.HZf[(v-18)*HZLD+i] : ((v>=30)?WQP.rho:real(1.0))*QQ(v,i);
            s+=z*hk; }
        col[i]=s; }
    WBC_SYNC();
    const real* rn=&RR(0,n);                                  // packed column n: RR(k,n) = rn[-k]
    for(int i=0;i<n;i++){
        real s=col[i]; const real* ri=&RR(0,i);
        for(int k=i-1;k>=0;k--) s-=ri[-k]*rn[-k];
        RR(i,n)=WBC_BYDIAG(s,w.rdR[i]);
    }
    real sum=col[n];
    for(int k=n-1;k>=0;k--) sum-=rn[-k]*rn[-k];
    if(!(sum>real(0.0))) return QP_RET_HESSIAN_NOT_SPD;
    const real d=sqrt(sum);
    RR(n,n)=d; w.rdR[n]=WBC_RDIAG(d);
    WBC_SYNC();
    return QP_OK;
}

// Data shift of the gradient and of the (far) bounds on x, QProblemB::determineDataShift (QProblemB.cpp:1694-1750): delta = target - current.  They are
// functions of the current g / lb / ub, the new gradient and the far bound, so they are re-evaluated where they are used instead of being kept in
// 1 KB of shared memory (what decides whether an SM holds 5 or 6 robots); g, lb and ub only change in the update of performStep, which reads them last.
template<class C> WBC_DEV real far_rv(const C& c,int idx);
WBC_DEV real g_target(const QPWork& w,int i){ return (i>=18&&i<30)?WQP.gf[i-18]:real(0.0); }
#define SHIFT_DG(i,gi)   (g_target(w,i)-(gi))        /* gi, lbi, ubi = the current g[i], lb[i], ub[i] (lane-owned vectors: loaded by the caller) */
#define SHIFT_DLB(i,lbi) (-far_rv(lv_h_,i)-(lbi))
#define SHIFT_DUB(i,ubi) (far_rv(lv_h_,i)-(ubi))

// ---------------------------------------------------------------- step direction
// QProblem::determineStepDirection (QProblem.cpp:4637-4975), HST_POSDEF branch.
// delta_* point to 42- / 70-long vectors by variable / constraint number.  Results: w.dx (free and fixed parts, by
// variable number), w.dyAC (by AC position), w.dyFX (by variable number).
// `dg` != null: the given gradient shift and no shift of bounds / constraint bounds (linear-independence tests); `dg` == null: the homotopy's own data shift
WBC_FN int determineStepDirection(QPWork& w,QPCtx& c,const real* dg,
                                  bool bCzero,bool bBzero){
    const int lane=WBC_LANE(c), nFR=c.nFR, nFX=c.nFX, nAC=c.nAC, nZ=nFR-nAC;
    if(!prepare_T(w,c)) return QP_RET_STEPDIRECTION_FAILED_TQ;
    const unsigned ffm=force_mask(w,c,0), fxm=force_mask(w,c,1);
    // I) delta_xFX
    // homotopy step: the gradient shift goes by variable number into w.tmpv (dead outside the working-set changes) so that both modes read dg[]; the
    // shift of the active constraints' bounds is in w.arow by AC position (determineDataShift)
    const bool shift=(dg==nullptr);
    if(shift){
        LV_FOR(v,sl,42){ LV_LD(gv,g,v,sl); LV_LD(lbv,lb,v,sl); LV_LD(ubv,ub,v,sl); LV_WAIT(); LV_USE(gv); LV_USE(lbv); LV_USE(ubv);
            if(LV_IN(v,42)){ const int bs=w.bstat[v]; const real dgv=SHIFT_DG(v,gv); w.tmpv[v]=dgv;
                if(bs!=0) w.dx[v]= bBzero?real(0.0):((bs==ST_LOWER)?SHIFT_DLB(v,lbv):SHIFT_DUB(v,ubv));
                else { w.tA[v]=dgv; w.dx[v]=real(0.0); } } }
        dg=w.tmpv;
    } else {
        LANE_FOR(i,nFX) w.dx[(int)w.FX[i]]=real(0.0);
        LANE_FOR(i,nFR){ int ii=(int)w.FR[i]; w.tA[ii]=dg[ii]; w.dx[ii]=real(0.0); }
    }
    LANE_FOR(i,nAC){ w.dyAC[i]=real(0.0); w.tB[i]= (bCzero||!shift)?real(0.0):w.arow[i]; }
    WBC_SYNC();
    int nref=0;
    for(int r=0;r<=OPT_numRefinementSteps;r++){
        if(nFR>0){
            LANE_FOR(i,nFR) w.xTMP[(int)w.FR[i]]=real(0.0);
            WBC_SYNC();
            if(nAC>0){
                if(bCzero&&bBzero){ LANE_FOR(i,nAC) w.xFRy[i]=real(0.0); WBC_SYNC(); }
                else {
                    if(!bBzero&&r==0){ mask_vars(w,c,w.dx,1); LANE_FOR(i,nAC) w.tB[i]=Arow_acc(WQP,(int)w.AC[i],w.vm,w.tB[i],-real(1.0)); WBC_SYNC(); }
                    backsolveT(w,c,w.tB,false,w.xFRy);
                    Q_times(w,c,nZ,nAC,w.xFRy,w.xTMP,nullptr);
                }
            }
            LANE_FOR(i,nZ) w.xFRz[i]=real(0.0);
            // HMX*delta_xFX and HFR*delta_xFRy into tempA
            if(!bBzero&&r==0){ LANE_FOR(i,nFR){ int ii=(int)w.FR[i]; w.tA[ii]=Hrow_acc(w,ii,w.dx,1,fxm,w.tA[ii],real(1.0)); } WBC_SYNC(); }
            if(nAC>0&&(!bCzero||!bBzero)){ LANE_FOR(i,nFR){ int ii=(int)w.FR[i]; w.tA[ii]=Hrow_acc(w,ii,w.xTMP,0,ffm,w.tA[ii],real(1.0)); } }
            WBC_SYNC();
            if(nZ>0){
                Qt_times(w,c,0,nZ,w.tA,-real(1.0),w.xFRz);
                backsolveR(w,c,w.xFRz,true,w.xFRz,nZ);
                backsolveR(w,c,w.xFRz,false,w.xFRz,nZ);
                Q_times(w,c,0,nZ,w.xFRz,w.xTMP,w.Zdz);
            }
        }
        // III) delta_yAC
        if(nAC>0){
            if(nZ>0){ LANE_FOR(i,nFR){ int ii=(int)w.FR[i]; w.tA[ii]=Hrow_acc(w,ii,w.Zdz,0,ffm,w.tA[ii],real(1.0)); } WBC_SYNC(); }
            Qt_times(w,c,nZ,nAC,w.tA,real(1.0),w.tB);
            backsolveT(w,c,w.tB,true,w.yTMP);
        }
        WBC_SYNC();
        LANE_FOR(i,nFR){ int ii=(int)w.FR[i]; w.dx[ii]+=w.xTMP[ii]; }
        LANE_FOR(i,nAC) w.dyAC[i]+=w.yTMP[i];
        WBC_SYNC();
        nref=r;
        if(r==OPT_numRefinementSteps) break;                 // the reference computes one more residual here and never reads it
        // residuals for the next refinement pass
        real rnrm=real(0.0);
        scatter_yAC(w,c,w.dyAC);
        LANE_FOR(i,nFR){ int ii=(int)w.FR[i]; real t=dg[ii]; t=Hrow_acc(w,ii,w.dx,0,ffm,t,real(1.0)); t=Hrow_acc(w,ii,w.dx,1,fxm,t,real(1.0)); t=Acol_acc(WQP,ii,w.ym,t,-real(1.0)); w.tA[ii]=t; rnrm=fmax(rnrm,fabs(t)); }
        // bA - A_FR dx_FR - A_FX dx_FX: the free columns first, then the fixed ones, each accumulated onto tempB (QProblem.cpp:4934-4935)
        mask_vars(w,c,w.dx,0);
        LANE_FOR(i,nAC){ int ii=(int)w.AC[i]; real t= (bCzero||!shift)?real(0.0):w.arow[i];
            w.tB[i]=Arow_acc(WQP,ii,w.vm,t,-real(1.0)); }
        if(!bBzero&&shift){ mask_vars(w,c,w.dx,1); LANE_FOR(i,nAC) w.tB[i]=Arow_acc(WQP,(int)w.AC[i],w.vm,w.tB[i],-real(1.0)); }
        LANE_FOR(i,nAC) rnrm=fmax(rnrm,fabs(w.tB[i]));
        rnrm=wbc_max(rnrm);
#if !defined(__CUDACC__) && defined(WBC_QP_DBG)
        if(g_dbg_iter==dbg_target()) printf("REF pass %d rnrm %a\n",r,rnrm);
#endif
        WBC_DBG_VEC("tmpxFR",w.xTMP[(int)w.FR[i]],nFR); WBC_DBG_VEC("xFRz",w.xFRz[i],nZ); WBC_DBG_VEC("xFRy",w.xFRy[i],nAC);
        WBC_DBG_VEC("resA",w.tA[(int)w.FR[i]],nFR); WBC_DBG_VEC("resB",w.tB[i],nAC);
        WBC_SYNC();
        if(rnrm<OPT_epsIterRef) break;
    }
    WBC_FLOPS((1.0+nref)*(2.0*nAC*nAC+4.0*nFR*nAC+4.0*nFR*nZ+4.0*nFR*nFR+4.0*nZ*nZ)+nref*(2.0*nFR*nFR+4.0*nAC*42.0)); (void)nref;
    // IV) delta_yFX
    scatter_yAC(w,c,w.dyAC);
    if(nFX>0){
        LV_FOR(v,sl,42){ real t=real(0.0);
            if(LV_IN(v,42)&&w.bstat[v]!=0){ t=dg[v]; t=Acol_acc(WQP,v,w.ym,t,-real(1.0)); t=Hrow_acc(w,v,w.dx,0,ffm,t,real(1.0)); t=Hrow_acc(w,v,w.dx,1,fxm,t,real(1.0)); }
            LV_ST(dyFX,v,sl,t); LV_DONE(); }
    }
    WBC_SYNC();
    WBC_DBG_VEC("dxFR",w.dx[(int)w.FR[i]],nFR); WBC_DBG_VEC("dyAC",w.dyAC[i],nAC); WBC_DBG_VEC("dxFX",w.dx[(int)w.FX[i]],nFX);
    return QP_OK;
}

// ---------------------------------------------------------------- ratio tests
// QProblemB::performRatioTest + isBlocking (QProblemB.cpp:2065-2107, QProblemB.ipp:474-488).  The reference scans the
// list sequentially, replacing the candidate whenever num_i < t*den_i with the CURRENT t.  Exactly the same chain of
// winners is produced here: a ballot finds the first later entry that beats the current t, t is updated, repeat.
// stat == 0: plain test (performPlainRatioTest, QProblem.cpp:4256-4275: num > epsNum, den > epsDen, t*den > num).
// num / den are indexed by the NUMBER of the variable / constraint (idx[i]); the scan runs in list order.
WBC_FN void performRatioTest(const QPCtx& c,int n,const signed char* idx,const signed char* stat,const real* num,const real* den,
                              real& t,int& BC_idx){
    const int lane=WBC_LANE(c); BC_idx=-1; int pos=0;
    WBC_SYNC();
    for(;;){
        int found=-1;
        for(int base=(pos/WBC_NL)*WBC_NL; base<n; base+=WBC_NL){
            int i=base+lane; bool f=false;
            if(i>=pos&&i<n){
                const int ii=(int)idx[i]; real nu=num[ii], de=den[ii];
                if(stat){ if(stat[ii]==ST_UPPER){ nu=-nu; de=-de; }
                          f=(de>=OPT_epsDen)&&(nu>=OPT_epsNum)&&(nu<t*de); }
                else f=(nu>OPT_epsNum)&&(de>OPT_epsDen)&&(t*de>nu);
            }
            unsigned m=wbc_ballot(f);
            if(m){ found=base+wbc_ffs(m); break; }
        }
        if(found<0) break;
        BC_idx=(int)idx[found]; t=num[BC_idx]/den[BC_idx]; pos=found+1;
    }
}

// ---------------------------------------------------------------- data shift, step, termination
// Targets of the current homotopy = the user's data with every infinite side replaced by a far bound (QProblem::updateFarBounds,
// QProblem.cpp:5498-5552, QProblemB.cpp:2202-2246: +-farbound*(1+ramp(i)), ramp offset as of the call).  There are no user bounds on
// x; the targets are functions of (index, farbound, offset) and are evaluated where needed instead of being stored.
template<class C> WBC_DEV real far_rv(const C& c,int idx){ const int nRamp=42+70; const real t=(real)((idx+c.farRampOffset)%nRamp)/(real)(nRamp-1);
    return c.farbound*(real(1.0)+(real(1.0)-t)*OPT_ramp0+t*OPT_ramp1); }
WBC_DEV real tgt_lb(const QPCtx& c,int i){ return -far_rv(c,i); }
WBC_DEV real tgt_ub(const QPCtx& c,int i){ return far_rv(c,i); }
template<class C> WBC_DEV real tgt_lbA(const QPWork& w,const C& c,int i){ return fmax(-far_rv(c,42+i),(real)c.qg->lbA[i]); }
template<class C> WBC_DEV real tgt_ubA(const QPWork& w,const C& c,int i){ return fmin(far_rv(c,42+i),(real)c.qg->ubA[i]); }
// QProblem::determineDataShift (QProblem.cpp:4571-4631) + QProblemB (:1694-1750); targets are WQP.gf (g_new) and the far-bounded data
WBC_FN void determineDataShift(QPWork& w,const QPCtx& c,bool& bCzero,bool& bBzero){
    const int lane=WBC_LANE(c); WBC_SYNC();
    bool nzC=false, nzB=false;
    // the shift of an ACTIVE constraint's bound is also left by AC position in w.arow (dead between two working-set changes): what determineStepDirection reads
    LV_FOR(i,sl,70){ LV_LD(lbAi,lbA,i,sl); LV_LD(ubAi,ubA,i,sl); LV_WAIT(); LV_USE(lbAi); LV_USE(ubAi);
        real dl=real(0.0), du=real(0.0);
        if(LV_IN(i,70)){ const real rv=far_rv(lv_h_,42+i); dl=fmax(-rv,(real)lv_h_.qg->lbA[i])-lbAi; du=fmin(rv,(real)lv_h_.qg->ubA[i])-ubAi;
            const int cs=w.cstat[i];
            if(cs!=0){ if(fabs(dl)>QEPS||fabs(du)>QEPS) nzC=true; w.arow[(int)w.posAC[i]]=(cs==ST_LOWER)?dl:du; } }
        LV_ST(dlbA,i,sl,dl); LV_ST(dubA,i,sl,du); LV_DONE(); }
    LV_FOR(i,sl,42){ LV_LD(lbi,lb,i,sl); LV_LD(ubi,ub,i,sl); LV_WAIT(); LV_USE(lbi); LV_USE(ubi);
        if(LV_IN(i,42)&&w.bstat[i]!=0){ if(fabs(SHIFT_DLB(i,lbi))>QEPS||fabs(SHIFT_DUB(i,ubi))>QEPS) nzB=true; } }
    WBC_SYNC();
    bBzero=(wbc_ballot(nzB)==0u);
    bCzero=(wbc_ballot(nzC)==0u);
}
// QProblem::getRelativeHomotopyLength (QProblem.cpp:5372-5410, QProblemB.cpp:2113-2157) is only ever compared with the termination
// tolerance: max_i |d_i|/s_i <= tol  <=>  no |d_i| > tol*s_i -- evaluated in that form (no fp64 divisions), s_i = max(|target_i|,1).
WBC_FN bool homotopyFinished(const QPWork& w,const QPCtx& c){
    const int lane=WBC_LANE(c); const real tol=OPT_terminationTolerance; bool over=false;
    LV_FOR(i,sl,42){ LV_LD(gi,g,i,sl); LV_LD(lbi,lb,i,sl); LV_LD(ubi,ub,i,sl); LV_WAIT(); LV_USE(gi); LV_USE(lbi); LV_USE(ubi);
        if(LV_IN(i,42)){ const real gn=g_target(w,i); if(fabs(gn-gi)>tol*fmax(fabs(gn),real(1.0))) over=true;
                         const real rv=far_rv(lv_h_,i), s=tol*fmax(rv,real(1.0)); if(fabs(-rv-lbi)>s||fabs(rv-ubi)>s) over=true; } }
    LV_FOR(i,sl,70){ LV_LD(lbAi,lbA,i,sl); LV_LD(ubAi,ubA,i,sl); LV_WAIT(); LV_USE(lbAi); LV_USE(ubAi);
        if(LV_IN(i,70)){ const real l=tgt_lbA(w,lv_h_,i), u=tgt_ubA(w,lv_h_,i); if(fabs(l-lbAi)>tol*fmax(fabs(l),real(1.0))||fabs(u-ubAi)>tol*fmax(fabs(u),real(1.0))) over=true; } }
    return wbc_ballot(over)==0u;
}
// QProblem::setupAuxiliaryQPgradient (QProblem.cpp:2602-2641): g = y_B - H x + A' y_C
WBC_FN void setupAuxiliaryQPgradient(QPWork& w,const QPCtx& c){
    const int lane=WBC_LANE(c); WBC_SYNC();
    LV_FOR(i,sl,42){ real t=real(0.0);
        if(LV_IN(i,42)){ t=w.y[i]; t=Hrow_acc(w,i,w.x,2,0xfffu,t,-real(1.0)); t=Acol_acc(WQP,i,w.y+42,t,real(1.0)); }
        LV_ST(g,i,sl,t); LV_DONE(); }
    WBC_SYNC();
}
// QProblem::performStep (QProblem.cpp:4981-5278)
WBC_FN void performStep(QPWork& w,QPCtx& c,int& BC_idx,int& BC_status,bool& BC_isBound){
    const int lane=WBC_LANE(c), nFR=c.nFR, nFX=c.nFX, nAC=c.nAC, nIAC=c.nIAC;
    c.tau=real(1.0); BC_idx=-1; BC_status=2 /*ST_UNDEFINED*/; BC_isBound=false; int tmp;
    WBC_FLOPS(2.0*nIAC*42.0+10.0*(42+70));
    WBC_SYNC();
    // I.1 active constraints' duals
    LANE_FOR(i,nAC){ int ii=(int)w.AC[i]; w.num[ii]=w.y[42+ii]; w.den[ii]=-w.dyAC[i]; }
    performRatioTest(c,nAC,w.AC,w.cstat,w.num,w.den,c.tau,tmp);
    if(tmp>=0){ BC_idx=tmp; BC_status=ST_INACTIVE; BC_isBound=false; }
    WBC_SYNC();
    // I.2 fixed variables' duals
    if(nFX>0){
        LV_FOR(v,sl,42){ LV_LD(dyv,dyFX,v,sl); LV_WAIT(); LV_USE(dyv); if(LV_IN(v,42)&&w.bstat[v]!=0){ w.num[v]=w.y[v]; w.den[v]=-dyv; } }
        performRatioTest(c,nFX,w.FX,w.bstat,w.num,w.den,c.tau,tmp);
        if(tmp>=0){ BC_idx=tmp; BC_status=ST_INACTIVE; BC_isBound=true; }
        WBC_SYNC();
    }
    // II.1 inactive constraints
    LANE_FOR(i,nIAC){ int ii=(int)w.IAC[i]; w.dAx[ii]=Arow_acc(WQP,ii,w.dx,real(0.0),real(1.0)); }
    WBC_SYNC();
    LV_FOR(i,sl,70){ LV_LD(al,Ax_l,i,sl); LV_LD(dl,dlbA,i,sl); LV_WAIT(); LV_USE(al); LV_USE(dl);
        if(LV_IN(i,70)&&w.cstat[i]==0){ w.num[i]=fmax(al,real(0.0)); w.den[i]=dl-w.dAx[i]; } }
    performRatioTest(c,nIAC,w.IAC,w.cstat,w.num,w.den,c.tau,tmp);
    if(tmp>=0){ BC_idx=tmp; BC_status=ST_LOWER; BC_isBound=false; }
    WBC_SYNC();
    LV_FOR(i,sl,70){ LV_LD(au,Ax_u,i,sl); LV_LD(du,dubA,i,sl); LV_WAIT(); LV_USE(au); LV_USE(du);
        if(LV_IN(i,70)&&w.cstat[i]==0){ w.num[i]=fmax(au,real(0.0)); w.den[i]=w.dAx[i]-du; } }
    performRatioTest(c,nIAC,w.IAC,w.cstat,w.num,w.den,c.tau,tmp);
    if(tmp>=0){ BC_idx=tmp; BC_status=ST_UPPER; BC_isBound=false; }
    WBC_SYNC();
    // II.2 free variables
    LV_FOR(v,sl,42){ LV_LD(lbv,lb,v,sl); LV_WAIT(); LV_USE(lbv);
        if(LV_IN(v,42)&&w.bstat[v]==0){ w.num[v]=fmax(w.x[v]-lbv,real(0.0)); w.den[v]=SHIFT_DLB(v,lbv)-w.dx[v]; } }
    performRatioTest(c,nFR,w.FR,w.bstat,w.num,w.den,c.tau,tmp);
    if(tmp>=0){ BC_idx=tmp; BC_status=ST_LOWER; BC_isBound=true; }
    WBC_SYNC();
    LV_FOR(v,sl,42){ LV_LD(ubv,ub,v,sl); LV_WAIT(); LV_USE(ubv);
        if(LV_IN(v,42)&&w.bstat[v]==0){ w.num[v]=fmax(ubv-w.x[v],real(0.0)); w.den[v]=w.dx[v]-SHIFT_DUB(v,ubv); } }
    performRatioTest(c,nFR,w.FR,w.bstat,w.num,w.den,c.tau,tmp);
    if(tmp>=0){ BC_idx=tmp; BC_status=ST_UPPER; BC_isBound=true; }
    WBC_SYNC();
    // III) step
    if(c.tau>QZERO){
        const real tau=c.tau;
        LANE_FOR(i,nFR){ int ii=(int)w.FR[i]; w.x[ii]+=tau*w.dx[ii]; }
        LANE_FOR(i,nAC){ int ii=(int)w.AC[i]; w.y[42+ii]+=tau*w.dyAC[i]; }
        LV_FOR(v,sl,42){ LV_LD(gv,g,v,sl); LV_LD(lbv,lb,v,sl); LV_LD(ubv,ub,v,sl); LV_LD(dyv,dyFX,v,sl); LV_WAIT(); LV_USE(gv); LV_USE(lbv); LV_USE(ubv); LV_USE(dyv);
            if(LV_IN(v,42)){ if(w.bstat[v]!=0){ w.x[v]+=tau*w.dx[v]; w.y[v]+=tau*dyv; }
                const real rv=far_rv(lv_h_,v); gv+=tau*SHIFT_DG(v,gv); lbv+=tau*(-rv-lbv); ubv+=tau*(rv-ubv); }
            LV_ST(g,v,sl,gv); LV_ST(lb,v,sl,lbv); LV_ST(ub,v,sl,ubv); LV_DONE(); }
        WBC_SYNC();
        // A x of the active constraints by number into w.num (the ratio tests are over), then one pass over the 70 constraints
        LANE_FOR(i,nAC){ int ii=(int)w.AC[i]; w.num[ii]=Arow_acc(WQP,ii,w.x,real(0.0),real(1.0)); }
        WBC_SYNC();
        LV_FOR(i,sl,70){ LV_LD(lbAi,lbA,i,sl); LV_LD(ubAi,ubA,i,sl); LV_LD(dl,dlbA,i,sl); LV_LD(du,dubA,i,sl); LV_LD(axi,Ax,i,sl); LV_LD(al,Ax_l,i,sl); LV_LD(au,Ax_u,i,sl);
            LV_WAIT(); LV_USE(lbAi); LV_USE(ubAi); LV_USE(dl); LV_USE(du); LV_USE(axi); LV_USE(al); LV_USE(au);
            if(LV_IN(i,70)){ lbAi+=tau*dl; ubAi+=tau*du;
                if(w.cstat[i]!=0){ const real ax=w.num[i]; axi=ax; au=ubAi-ax; al=ax-lbAi; }
                else { const real da=w.dAx[i]; axi+=tau*da; al+=tau*(da-dl); au+=tau*(du-da); } }
            LV_ST(lbA,i,sl,lbAi); LV_ST(ubA,i,sl,ubAi); LV_ST(Ax,i,sl,axi); LV_ST(Ax_l,i,sl,al); LV_ST(Ax_u,i,sl,au); LV_DONE(); }
        WBC_SYNC();
    }
}
// QProblem::performRamping (QProblem.cpp:5416-5492); all bounds and constraints are typed ST_BOUNDED on this path
WBC_FN void performRamping(QPWork& w,QPCtx& c){
    const int lane=WBC_LANE(c), nRamp=42+70+70+42; WBC_SYNC();
    LV_FOR(i,sl,42){ LV_LD(lbi,lb,i,sl); LV_LD(ubi,ub,i,sl); LV_WAIT(); LV_USE(lbi); LV_USE(ubi);
        if(LV_IN(i,42)){
            real tP=(real)((i+c.rampOffset)%nRamp)/(real)(nRamp-1), rP=(real(1.0)-tP)*OPT_ramp0+tP*OPT_ramp1;
            real tD=(real)((42+70+70+i+c.rampOffset)%nRamp)/(real)(nRamp-1), rD=(real(1.0)-tD)*OPT_ramp0+tD*OPT_ramp1;
            int bs=w.bstat[i]; real xi=w.x[i], sca=fmax(fabs(xi),real(1.0));
            if(bs!=ST_LOWER) lbi=xi-sca*rP;
            if(bs!=ST_UPPER) ubi=xi+sca*rP;
            if(bs==ST_LOWER){ lbi=xi; w.y[i]=+rD; }
            if(bs==ST_UPPER){ ubi=xi; w.y[i]=-rD; }
            if(bs==ST_INACTIVE) w.y[i]=real(0.0);
        }
        LV_ST(lb,i,sl,lbi); LV_ST(ub,i,sl,ubi); LV_DONE(); }
    LV_FOR(i,sl,70){ LV_LD(lbAi,lbA,i,sl); LV_LD(ubAi,ubA,i,sl); LV_LD(ax,Ax,i,sl); LV_WAIT(); LV_USE(lbAi); LV_USE(ubAi); LV_USE(ax);
        real al=real(0.0), au=real(0.0);
        if(LV_IN(i,70)){
            real tP=(real)((42+i+c.rampOffset)%nRamp)/(real)(nRamp-1), rP=(real(1.0)-tP)*OPT_ramp0+tP*OPT_ramp1;
            real tD=(real)((42+70+i+c.rampOffset)%nRamp)/(real)(nRamp-1), rD=(real(1.0)-tD)*OPT_ramp0+tD*OPT_ramp1;
            int cs=w.cstat[i]; real sca=fmax(fabs(ax),real(1.0));
            if(cs!=ST_LOWER) lbAi=ax-sca*rP;
            if(cs!=ST_UPPER) ubAi=ax+sca*rP;
            if(cs==ST_LOWER){ lbAi=ax; w.y[42+i]=+rD; }
            if(cs==ST_UPPER){ ubAi=ax; w.y[42+i]=-rD; }
            if(cs==ST_INACTIVE) w.y[42+i]=real(0.0);
            al=ax-lbAi; au=ubAi-ax;
        }
        LV_ST(lbA,i,sl,lbAi); LV_ST(ubA,i,sl,ubAi); LV_ST(Ax_l,i,sl,al); LV_ST(Ax_u,i,sl,au); LV_DONE(); }
    setupAuxiliaryQPgradient(w,c);
    c.rampOffset++;
}
// QProblem::performDriftCorrection (QProblem.cpp:5559-5652)
WBC_FN void performDriftCorrection(QPWork& w,QPCtx& c){
    const int lane=WBC_LANE(c); WBC_SYNC();
    LV_FOR(i,sl,42){ LV_LD(lbi,lb,i,sl); LV_LD(ubi,ub,i,sl); LV_WAIT(); LV_USE(lbi); LV_USE(ubi);
        if(LV_IN(i,42)){ int bs=w.bstat[i]; real xi=w.x[i];
            if(bs==ST_LOWER){ lbi=xi; ubi=fmax(ubi,xi); w.y[i]=fmax(w.y[i],real(0.0)); }
            else if(bs==ST_UPPER){ lbi=fmin(lbi,xi); ubi=xi; w.y[i]=fmin(w.y[i],real(0.0)); }
            else { lbi=fmin(lbi,xi); ubi=fmax(ubi,xi); w.y[i]=real(0.0); } }
        LV_ST(lb,i,sl,lbi); LV_ST(ub,i,sl,ubi); LV_DONE(); }
    LV_FOR(i,sl,70){ LV_LD(lbAi,lbA,i,sl); LV_LD(ubAi,ubA,i,sl); LV_LD(ax,Ax,i,sl); LV_WAIT(); LV_USE(lbAi); LV_USE(ubAi); LV_USE(ax);
        real al=real(0.0), au=real(0.0);
        if(LV_IN(i,70)){ int cs=w.cstat[i];
            if(cs==ST_LOWER){ lbAi=ax; al=real(0.0); ubAi=fmax(ubAi,ax); au=ubAi-ax; w.y[42+i]=fmax(w.y[42+i],real(0.0)); }
            else if(cs==ST_UPPER){ lbAi=fmin(lbAi,ax); al=ax-lbAi; ubAi=ax; au=real(0.0); w.y[42+i]=fmin(w.y[42+i],real(0.0)); }
            else { lbAi=fmin(lbAi,ax); al=ax-lbAi; ubAi=fmax(ubAi,ax); au=ubAi-ax; w.y[42+i]=real(0.0); } }
        LV_ST(lbA,i,sl,lbAi); LV_ST(ubA,i,sl,ubAi); LV_ST(Ax_l,i,sl,al); LV_ST(Ax_u,i,sl,au); LV_DONE(); }
    setupAuxiliaryQPgradient(w,c);
}

// ---------------------------------------------------------------- Givens (QProblemB.ipp:409-470, four-multiplication form)
WBC_DEV void computeGivens(real xold,real yold,real& xnew,real& ynew,real& cc,real& ss){
    if(fabs(yold)<QZERO){ cc=real(1.0); ss=real(0.0); xnew=xold; ynew=yold; }
    else { real mu=fabs(xold); if(fabs(yold)>mu) mu=fabs(yold);
           real t=mu*sqrt((xold/mu)*(xold/mu)+(yold/mu)*(yold/mu)); if(xold<real(0.0)) t=-t;
           cc=xold/t; ss=yold/t; xnew=t; ynew=real(0.0); }
}
// The column-wise Givens chains of addConstraint / addBound: rotation j is defined by (wv[j+1], wv[j]) where wv[j] is the
// result of rotation j-1 -- a scalar recurrence every lane evaluates redundantly -- and then applied to the column
// pair (j+1, j) of Q.  The rows of Q are independent, so each lane sweeps its own rows through the whole chain.
WBC_FN void givens_chain(QPWork& w,const QPCtx& c,int nrot){
    // The recurrence t_j = hypot(w[j+1], t_{j-1}) of the reference's loop is t_j = sgn(w[j+1]) |w[0..j+1]|_2, so every
    // rotation follows from prefix sums of squares: one lane per rotation, no serial sqrt/divide chain.
    const int lane=WBC_LANE(c); WBC_SYNC();
#if defined(WBC_EXACT_GIVENS)
    { // the reference's serial recurrence (QProblem.cpp:2909-2925, 3401-3450): rotation j is computed from (wv[j+1], result of rotation j-1); every lane
      // evaluates it redundantly, lane 0 stores
        real prev=(nrot>0)?w.wv[0]:real(0.0);
        for(int j=0;j<nrot;j++){ real xn,yn,cc,ss; computeGivens(w.wv[j+1],prev,xn,yn,cc,ss); if(lane==0){ w.cs_c[j]=cc; w.cs_s[j]=ss; } prev=xn; }
        WBC_SYNC();
        if(nrot>0&&lane==0) w.wv[nrot]=prev;
        WBC_SYNC();
        return;
    }
#endif
    real tlast=w.wv[0];
    LANE_FOR(j,nrot){
        real P=real(0.0); for(int k=0;k<=j;k++) P+=w.wv[k]*w.wv[k];      // sequential on purpose: a warp scan re-associates the sum and moved one degenerate working set
        const real xold=w.wv[j+1];
        const real yold=(j==0)?w.wv[0]:((w.wv[j]<real(0.0))?-sqrt(P):sqrt(P));
        real cc,ss,t;
        if(fabs(yold)<QZERO){ cc=real(1.0); ss=real(0.0); t=xold; }
        else { t=sqrt(P+xold*xold); if(xold<real(0.0)) t=-t; const real rt=real(1.0)/t; cc=xold*rt; ss=yold*rt; }
        w.cs_c[j]=cc; w.cs_s[j]=ss; if(j==nrot-1) tlast=t;
    }
    WBC_SYNC();
    if(nrot>0){ const int owner=(nrot-1)%WBC_NL; tlast=wbc_bcast(tlast,owner); if(lane==0) w.wv[nrot]=tlast; }
    WBC_SYNC();
}
WBC_FN void apply_chain_Q(QPWork& w,const QPCtx& c,int nrot,int nrows){
    const int lane=WBC_LANE(c);
#if WBC_NL==1
    for(int i=0;i<nrows;i++){ int ii=(int)w.FR[i]; real qa=QQ(ii,0);
        for(int j=0;j<nrot;j++){ real qb=QQ(ii,j+1), cc=w.cs_c[j], ss=w.cs_s[j]; QQ(ii,j)=-ss*qb+cc*qa; qa=cc*qb+ss*qa; }
        QQ(ii,nrot)=qa; }
#else
    if(nrows>32){
        // more rows than lanes: lane l < ceil(nrows/2) sweeps rows l and l+half together (two independent chains, the rotation is
        // loaded once, no predicates in the loop); the other lanes have nothing to do
        const int half=(nrows+1)>>1;
        if(lane<half){
            const bool two=lane+half<nrows;
            real* r0=&w.Q[(int)w.FR[lane]]; real* r1=&w.Q[(int)w.FR[two?lane+half:lane]];
            if(two){
                real a0=r0[0], a1=r1[0];
                #pragma unroll 2
                for(int j=0;j<nrot;j++){ const real cc=w.cs_c[j], ss=w.cs_s[j], b0=r0[LD*(j+1)], b1=r1[LD*(j+1)];
                    r0[LD*j]=-ss*b0+cc*a0; r1[LD*j]=-ss*b1+cc*a1; a0=cc*b0+ss*a0; a1=cc*b1+ss*a1; }
                r0[LD*nrot]=a0; r1[LD*nrot]=a1;
            } else {                                             // odd row count: the last lane of the first half has one row
                real a0=r0[0];
                for(int j=0;j<nrot;j++){ const real cc=w.cs_c[j], ss=w.cs_s[j], b0=r0[LD*(j+1)]; r0[LD*j]=-ss*b0+cc*a0; a0=cc*b0+ss*a0; }
                r0[LD*nrot]=a0;
            }
        }
    } else if(lane<nrows){
        real* r0=&w.Q[(int)w.FR[lane]];
        real a0=r0[0];
        #pragma unroll 2
        for(int j=0;j<nrot;j++){ const real cc=w.cs_c[j], ss=w.cs_s[j], b0=r0[LD*(j+1)]; r0[LD*j]=-ss*b0+cc*a0; a0=cc*b0+ss*a0; }
        r0[LD*nrot]=a0;
    }
#endif
    WBC_SYNC();
}

WBC_FN int removeBound(QPWork& w,QPCtx& c,int number);
WBC_FN int removeConstraint(QPWork& w,QPCtx& c,int number);

// QProblem::addConstraint_checkLI (QProblem.cpp:2982-3106): returns true if linearly independent
WBC_FN bool addConstraint_checkLI(QPWork& w,QPCtx& c,int number,bool fullTest){
    const int lane=WBC_LANE(c), nFR=c.nFR, nZ=c.nFR-c.nAC;
    WBC_SYNC();
    if(fullTest){
        int shortcut=-1; (void)shortcut;
        LANE_FOR(j,42) w.tmpv[j]=Ael_dev(WQP,number,j);
        WBC_SYNC();
        { // The full test costs a whole step-direction solve and only its verdict is used.  |Z'a| decides it beforehand in all but
          // borderline cases: the full test's "zero" is |Z (Z'HZ)^-1 Z'a|, which vanishes with Z'a (measured on 6655 tests of
          // this QP family: independent rows have max|Z'a| >= 1e-2 |a|, dependent ones exactly 0).  Only the band in between
          // runs the reference's test.
            real l2=real(0.0); for(int i=0;i<nFR;i++){ const real a=w.tmpv[(int)w.FR[i]]; l2+=a*a; }
            Qt_times(w,c,0,nZ,w.tmpv,real(1.0),w.xFRz); real mx=real(0.0); LANE_FOR(j,nZ) mx=fmax(mx,fabs(w.xFRz[j]));
            mx=wbc_max(mx);
            const real nrm=sqrt(l2);
            const bool sureLI=mx>LI_SURE*nrm, sureLD=mx<=LI_NONE*nrm;
#if defined(WBC_EXACT)
            shortcut= sureLI?1:(sureLD?0:-1);
#else
            if(sureLI||sureLD){ WBC_FLOPS(3.0*(2.0*c.nAC*c.nAC+4.0*nFR*c.nAC+4.0*nFR*nZ+4.0*nFR*nFR+4.0*nZ*nZ)+2.0*(2.0*nFR*nFR+4.0*c.nAC*42.0)); return sureLI; }
#endif
        }
        int rv=determineStepDirection(w,c,w.tmpv,false,false);
        if(rv!=QP_OK) return false;
        real weight=real(0.0), zero=real(0.0);
        LANE_FOR(i,c.nAC) weight=fmax(weight,fabs(w.dyAC[i]));
        if(c.nFX>0){ LV_FOR(v,sl,42){ LV_LD(dyv,dyFX,v,sl); LV_WAIT(); LV_USE(dyv); if(LV_IN(v,42)&&w.bstat[v]!=0){ weight=fmax(weight,fabs(dyv)); zero=fmax(zero,fabs(w.dx[v])); } } }
        LANE_FOR(i,nFR) zero=fmax(zero,fabs(w.dx[(int)w.FR[i]]));
        weight=wbc_max(weight); zero=wbc_max(zero);
#if !defined(__CUDACC__) && defined(WBC_LI_STATS)
        { real l2=real(0.0),mx=real(0.0),amax=real(0.0); for(int i=0;i<nFR;i++){ real a=w.tmpv[(int)w.FR[i]]; l2+=a*a; amax=fmax(amax,fabs(a)); }
          for(int j=0;j<nZ;j++){ real sum=real(0.0); for(int i=0;i<nFR;i++){ int ii=(int)w.FR[i]; sum+=w.tmpv[ii]*QQ(ii,j); } mx=fmax(mx,fabs(sum)); }
          printf("LISTAT C %d %.3e %.3e %.3e %.3e %.3e\n",(int)(zero>OPT_epsLITests*weight),mx,l2,amax,zero,weight); }
#endif
#if defined(WBC_EXACT)
        { const bool full=zero>OPT_epsLITests*weight; WBC_LI_COUNT(0); if(shortcut<0) WBC_LI_COUNT(1) else if((shortcut==1)!=full) WBC_LI_COUNT(2) }
#endif
        return zero>OPT_epsLITests*weight;
    }
    LANE_FOR(j,42) w.arow[j]=Ael_dev(WQP,number,j);
    WBC_SYNC();
    real l2=real(0.0); for(int i=0;i<nFR;i++){ real a=w.arow[(int)w.FR[i]]; l2+=a*a; }
    bool li=false;
    Qt_times(w,c,0,nZ,w.arow,real(1.0),w.xFRz); LANE_FOR(j,nZ) if(fabs(w.xFRz[j])>OPT_epsLITests*l2) li=true;
    return wbc_ballot(li)!=0u;
}
// QProblem::addBound_checkLI (QProblem.cpp:3487-3590)
WBC_FN bool addBound_checkLI(QPWork& w,QPCtx& c,int number,bool fullTest){
    const int lane=WBC_LANE(c), nZ=c.nFR-c.nAC;
    WBC_SYNC();
    if(fullTest){
        int shortcut=-1; (void)shortcut;
        { // same shortcut as addConstraint_checkLI with a = e_number
            real mx=real(0.0); LANE_FOR(i,nZ) mx=fmax(mx,fabs(QQ(number,i)));
            mx=wbc_max(mx);
            const bool sureLI=mx>LI_SURE, sureLD=mx<=LI_NONE;
#if defined(WBC_EXACT)
            shortcut= sureLI?1:(sureLD?0:-1);
#else
            if(sureLI||sureLD){ WBC_FLOPS(3.0*(2.0*c.nAC*c.nAC+4.0*c.nFR*c.nAC+4.0*c.nFR*nZ+4.0*c.nFR*c.nFR+4.0*nZ*nZ)+2.0*(2.0*c.nFR*c.nFR+4.0*c.nAC*42.0)); return sureLI; }
#endif
        }
        LANE_FOR(j,42) w.tmpv[j]=(j==number)?real(1.0):real(0.0);
        WBC_SYNC();
        int rv=determineStepDirection(w,c,w.tmpv,false,false);
        if(rv!=QP_OK) return false;
        real weight=real(0.0), zero=real(0.0);
        LANE_FOR(i,c.nAC) weight=fmax(weight,fabs(w.dyAC[i]));
        if(c.nFX>0){ LV_FOR(v,sl,42){ LV_LD(dyv,dyFX,v,sl); LV_WAIT(); LV_USE(dyv); if(LV_IN(v,42)&&w.bstat[v]!=0){ weight=fmax(weight,fabs(dyv)); zero=fmax(zero,fabs(w.dx[v])); } } }
        LANE_FOR(i,c.nFR) zero=fmax(zero,fabs(w.dx[(int)w.FR[i]]));
        weight=wbc_max(weight); zero=wbc_max(zero);
#if !defined(__CUDACC__) && defined(WBC_LI_STATS)
        { double mx=0.0; for(int i=0;i<nZ;i++) mx=fmax(mx,fabs(QQ(number,i))); printf("LISTAT B %d %.3e 1 1 %.3e %.3e\n",(int)(zero>OPT_epsLITests*weight),mx,zero,weight); }
#endif
#if defined(WBC_EXACT)
        { const bool full=zero>OPT_epsLITests*weight; WBC_LI_COUNT(0); if(shortcut<0) WBC_LI_COUNT(1) else if((shortcut==1)!=full) WBC_LI_COUNT(2) }
#endif
        return zero>OPT_epsLITests*weight;
    }
    bool li=false; LANE_FOR(i,nZ) if(fabs(QQ(number,i))>OPT_epsLITests) li=true;
    return wbc_ballot(li)!=0u;
}

// common tail of addConstraint_ensureLI / addBound_ensureLI (QProblem.cpp:3178-3300, 3658-3775): xiC in w.xFRy (AC position),
// xiB in w.Zdz (variable number).  returns QP_OK, or QP_RET_HOTSTART_STOPPED_INFEASIBILITY if no index can leave.
WBC_FN int ensureLI_exchange(QPWork& w,QPCtx& c,int number,int status,bool isBound){
    const int lane=WBC_LANE(c), nAC=c.nAC, nFX=c.nFX;
    real y_min=OPT_maxDualJump; int ymn=-1, ymnb=-1;
    WBC_SYNC();
    LANE_FOR(i,nAC){ int ii=(int)w.AC[i]; w.num[ii]=w.y[42+ii]; w.den[ii]=w.xFRy[i]; }
    performRatioTest(c,nAC,w.AC,w.cstat,w.num,w.den,y_min,ymn);
    WBC_SYNC();
    LANE_FOR(i,nFX){ int ii=(int)w.FX[i]; w.num[ii]=w.y[ii]; w.den[ii]=w.Zdz[ii]; }
    performRatioTest(c,nFX,w.FX,w.bstat,w.num,w.den,y_min,ymnb);
    bool ymIsBound=false; if(ymnb>=0){ ymn=ymnb; ymIsBound=true; }
    if(ymn<0){ c.infeasible=1; return QP_RET_HOTSTART_STOPPED_INFEASIBILITY; }
    WBC_SYNC();
    LANE_FOR(i,nAC){ int ii=(int)w.AC[i]; w.y[42+ii]-=y_min*w.xFRy[i]; }
    LANE_FOR(i,nFX){ int ii=(int)w.FX[i]; w.y[ii]-=y_min*w.Zdz[ii]; }
    WBC_SYNC();
    if(lane==0){ real v=(status==ST_LOWER)?y_min:-y_min; if(isBound) w.y[number]=v; else w.y[42+number]=v; }
    WBC_SYNC();
    int rv;
    if(ymIsBound){ rv=removeBound(w,c,ymn); if(rv!=QP_OK) return rv; if(lane==0) w.y[ymn]=real(0.0); }
    else { rv=removeConstraint(w,c,ymn); if(rv!=QP_OK) return rv; if(lane==0) w.y[42+ymn]=real(0.0); }
    WBC_SYNC();
    return QP_OK;
}
// QProblem::addConstraint_ensureLI (QProblem.cpp:3117-3300)
WBC_FN int addConstraint_ensureLI(QPWork& w,QPCtx& c,int number,int C_status){
    const int lane=WBC_LANE(c);
    if(addConstraint_checkLI(w,c,number,true)) return QP_OK;
    const int nFR=c.nFR, nAC=c.nAC, nFX=c.nFX, nZ=nFR-nAC; const real sg=(C_status==ST_LOWER)?real(1.0):-real(1.0);
    WBC_SYNC();
    if(nAC>0){
        LANE_FOR(j,42) w.arow[j]=sg*Ael_dev(WQP,number,j);
        WBC_SYNC();
        Qt_times(w,c,nZ,nAC,w.arow,real(1.0),w.tB);
        if(!prepare_T(w,c)) return QP_RET_UNSUPPORTED_PATH;
        backsolveT(w,c,w.tB,true,w.xFRy);
    }
    WBC_SYNC();
    scatter_yAC(w,c,w.xFRy);
    LANE_FOR(i,nFX){ int ii=(int)w.FX[i]; w.Zdz[ii]=Acol_acc(WQP,ii,w.ym,sg*Ael_dev(WQP,number,ii),-real(1.0)); }
    return ensureLI_exchange(w,c,number,C_status,false);
}
// QProblem::addBound_ensureLI (QProblem.cpp:3596-3780)
WBC_FN int addBound_ensureLI(QPWork& w,QPCtx& c,int number,int B_status){
    const int lane=WBC_LANE(c);
    if(addBound_checkLI(w,c,number,true)) return QP_OK;
    const int nAC=c.nAC, nFX=c.nFX, nZ=c.nFR-nAC; const real sg=(B_status==ST_LOWER)?real(1.0):-real(1.0);
    WBC_SYNC();
    if(nAC>0){
        LANE_FOR(i,nAC) w.tB[i]=sg*QQ(number,nZ+i);
        if(!prepare_T(w,c)) return QP_RET_UNSUPPORTED_PATH;
        backsolveT(w,c,w.tB,true,w.xFRy);
    }
    WBC_SYNC();
    scatter_yAC(w,c,w.xFRy);
    LANE_FOR(i,nFX){ int ii=(int)w.FX[i]; w.Zdz[ii]=Acol_acc(WQP,ii,w.ym,real(0.0),-real(1.0)); }
    return ensureLI_exchange(w,c,number,B_status,true);
}

// QProblem::addConstraint (QProblem.cpp:2819-2975), second half: w.wv = aFR' [Z Y] is given; new row of T, Givens chain on Z, index lists
WBC_FN void addConstraint_tail(QPWork& w,QPCtx& c,int number,int C_status){
    const int lane=WBC_LANE(c), nFR=c.nFR, nAC=c.nAC, nZ=nFR-nAC, tcol=SIZET-nAC;
    WBC_FLOPS(6.0*nFR*nZ+3.0*nZ*nZ+2.0*nFR*42.0);
    LANE_FOR(j,nAC) TT(nAC,tcol+j)=w.wv[nZ+j];
    WBC_SYNC();
    if(nZ>0){
        givens_chain(w,c,nZ-1);
        apply_chain_Q(w,c,nZ-1,nFR);
        if(lane==0) TT(nAC,tcol-1)=(nZ>1)?w.wv[nZ-1]:w.wv[0];
    }
    moveInactiveToActive(w,c,number,C_status);
}
WBC_FN int addConstraint(QPWork& w,QPCtx& c,int number,int C_status,bool updateCholesky,bool ensureLI){
    const int lane=WBC_LANE(c);
    if(updateCholesky&&ensureLI){ int rv=addConstraint_ensureLI(w,c,number,C_status); if(rv!=QP_OK) return rv; }
    WBC_SYNC();
    // row [wZ wY] = aFR' [Z Y]
    LANE_FOR(j,42) w.arow[j]=Ael_dev(WQP,number,j);
    WBC_SYNC();
    Qt_times(w,c,0,c.nFR,w.arow,real(1.0),w.wv);
    addConstraint_tail(w,c,number,C_status);
    return QP_OK;
}
// setupAuxiliaryWorkingSet step of the warm start (SQProblem.cpp:340-515 -> QProblem.cpp:2212-2378): cheap LI test
// (addConstraint_checkLI, QProblem.cpp:3081-3105) and addConstraint without LI exchange.  Both need aFR'Z; it is computed once
// (the first nZ entries of aFR'[Z Y] are the very same sums).
template<class QD> WBC_FN void setup_readd_constraint(QPWork& w,QPCtx& c,const QD& qd,int number,int C_status){
    const int lane=WBC_LANE(c), nFR=c.nFR, nZ=c.nFR-c.nAC;
    WBC_SYNC();
    LANE_FOR(j,42) w.arow[j]=Ael_dev(qd,number,j);
    WBC_SYNC();
    real l2=real(0.0);
    if(c.frIdentity){
        #pragma unroll 6
        for(int i=0;i<42;i++){ const real a=w.arow[i]; l2+=a*a; }
    } else for(int i=0;i<nFR;i++){ const real a=w.arow[(int)w.FR[i]]; l2+=a*a; }
    Qt_times(w,c,0,nFR,w.arow,real(1.0),w.wv);
    bool li=false; LANE_FOR(j,nZ) if(fabs(w.wv[j])>OPT_epsLITests*l2) li=true;
    if(wbc_ballot(li)==0u) return;
    addConstraint_tail(w,c,number,C_status);
}
// QProblem::addBound (QProblem.cpp:3307-3480)
WBC_FN int addBound(QPWork& w,QPCtx& c,int number,int B_status,bool updateCholesky,bool ensureLI){
    const int lane=WBC_LANE(c);
    if(updateCholesky&&ensureLI){ int rv=addBound_ensureLI(w,c,number,B_status); if(rv!=QP_OK) return rv; }
    const int nFR=c.nFR, nAC=c.nAC, nZ=nFR-nAC, tcol=SIZET-nAC;
    WBC_FLOPS(6.0*nFR*nZ+3.0*nZ*nZ+2.0*nFR*42.0);
#if !defined(__CUDACC__) && defined(WBC_QP_TRACE)
    printf("    addBound %2d  nFR %2d nAC %2d\n",number,nFR,nAC);
#endif
    WBC_SYNC();
    const int lastfree=(int)w.FR[nFR-1];
    if(lastfree!=number) swapFree(w,c,number,lastfree);
    LANE_FOR(i,nFR) w.wv[i]=QQ(number,i);
    const int nrot=nFR-1;                          // j = 0 .. nFR-2 (the three loops of the reference are contiguous)
    givens_chain(w,c,(nAC>0)?nrot:(nZ-1));
    apply_chain_Q(w,c,(nAC>0)?nrot:(nZ-1),nFR);
    if(nAC>0){
        // T rows: rotation j >= nZ-1 touches rows i >= nFR-2-j at column 1+tcol-nZ+j, paired with the temporary column tmp
        LANE_FOR(i,nAC){ real tmp=real(0.0); int j0=nFR-2-i; if(j0<nZ-1) j0=nZ-1;
            for(int j=j0;j<=nFR-2;j++){ int col=1+tcol-nZ+j; real xo=TT(i,col), cc=w.cs_c[j], ss=w.cs_s[j];
                real xn=cc*xo+ss*tmp, yn=-ss*xo+cc*tmp; tmp=xn; TT(i,col)=yn; } }
    }
    moveFreeToFixed(w,c,number,B_status);
    return QP_OK;
}
// QProblem::removeConstraint (QProblem.cpp:3782-4012) without NZC test / flipper / R update (see file header)
WBC_FN int removeConstraint(QPWork& w,QPCtx& c,int number){
    const int lane=WBC_LANE(c), nFR=c.nFR, nAC=c.nAC, nZ=nFR-nAC, tcol=SIZET-nAC;
    WBC_FLOPS(6.0*nFR*nZ+3.0*nZ*nZ+2.0*nFR*42.0);
    WBC_SYNC();
    const int number_idx=(int)w.posAC[number];
    if(number_idx<nAC-1){
        // shift rows up; lane per reverse-column m so that the upward copy is ordered inside one lane.  Row i has i+1 entries
        // (m = 0..i); moved to row i-1 it keeps m = 0..i-1 in place and its leading entry m = i -- the one the Givens sweep
        // below annihilates -- in w.wv[i-1] (packed storage has no room for it).
        LANE_FOR(m,nAC){ for(int i=number_idx+1;i<nAC;i++){ if(m<i) TT(i-1,SIZET-1-m)=TT(i,SIZET-1-m); else if(m==i) w.wv[i-1]=TT(i,SIZET-1-m); } }
        for(int j=nAC-2-number_idx;j>=0;j--){
            WBC_SYNC();
            const int pr=nAC-2-j; real xn,yn,cc,ss; computeGivens(TT(pr,tcol+1+j),w.wv[pr],xn,yn,cc,ss);
            WBC_SYNC();
            if(lane==0){ TT(pr,tcol+1+j)=xn; w.wv[pr]=yn; }
            for(int i=nAC-j-1+lane;i<nAC-1;i+=WBC_NL){ real xo=TT(i,tcol+1+j), yo=TT(i,tcol+j); TT(i,tcol+1+j)=cc*xo+ss*yo; TT(i,tcol+j)=-ss*xo+cc*yo; }
            LANE_FOR(i,nFR){ int ii=(int)w.FR[i]; real xo=QQ(ii,nZ+1+j), yo=QQ(ii,nZ+j); QQ(ii,nZ+1+j)=cc*xo+ss*yo; QQ(ii,nZ+j)=-ss*xo+cc*yo; }
        }
    }
    moveActiveToInactive(w,c,number);
    return QP_OK;
}
// QProblem::removeBound (QProblem.cpp:4018-4252) without NZC test / flipper / R update (see file header)
WBC_FN int removeBound(QPWork& w,QPCtx& c,int number){
    const int lane=WBC_LANE(c), nFR=c.nFR, nAC=c.nAC, nZ=nFR-nAC, tcol=SIZET-nAC;
    WBC_FLOPS(6.0*nFR*nZ+3.0*nZ*nZ+2.0*nFR*42.0);
    moveFixedToFree(w,c,number);
    LANE_FOR(i,nFR){ int ii=(int)w.FR[i]; QQ(ii,nFR)=real(0.0); QQ(number,i)=real(0.0); }
    WBC_SYNC();
    if(lane==0) QQ(number,nFR)=real(1.0);
    if(nAC>0){
        LANE_FOR(i,nAC) w.tmpv[i]=Ael_dev(WQP,(int)w.AC[i],number);
        for(int j=nAC-1;j>=0;j--){
            WBC_SYNC();
            const int pr=nAC-1-j; real xn,yn,cc,ss; computeGivens(w.tmpv[pr],TT(pr,tcol+j),xn,yn,cc,ss);
            WBC_SYNC();
            if(lane==0){ TT(pr,tcol+j)=xn; w.tmpv[pr]=yn; }
            for(int i=nAC-j+lane;i<nAC;i+=WBC_NL){ real xo=w.tmpv[i], yo=TT(i,tcol+j); TT(i,tcol+j)=cc*xo+ss*yo; w.tmpv[i]=-ss*xo+cc*yo; }
            LANE_FOR(i,nFR+1){ int ii=(int)w.FR[i]; real xo=QQ(ii,nZ+1+j), yo=QQ(ii,nZ+j); QQ(ii,nZ+1+j)=cc*xo+ss*yo; QQ(ii,nZ+j)=-ss*xo+cc*yo; }
        }
    }
    WBC_SYNC();
    return QP_OK;
}

// ---------------------------------------------------------------- homotopy
// QProblem::changeActiveSet (QProblem.cpp:5284-5365)
WBC_FN int changeActiveSet(QPWork& w,QPCtx& c,int BC_idx,int BC_status,bool BC_isBound){
    int rv=QP_OK;
    if(BC_status==2) return QP_OK;
    if(BC_status==ST_INACTIVE){
        if(BC_isBound){ rv=removeBound(w,c,BC_idx); if(rv!=QP_OK) return rv; if(WBC_LANE(c)==0) w.y[BC_idx]=real(0.0); }
        else { rv=removeConstraint(w,c,BC_idx); if(rv!=QP_OK) return rv; if(WBC_LANE(c)==0) w.y[42+BC_idx]=real(0.0); }
        WBC_SYNC();
    } else {
        if(BC_isBound) rv=addBound(w,c,BC_idx,BC_status,true,true);
        else rv=addConstraint(w,c,BC_idx,BC_status,true,true);
    }
    return rv;
}
// QProblem::updateFarBounds (QProblem.cpp:5498-5552, QProblemB.cpp:2202-2246): see far_rv
WBC_DEV void updateFarBounds(QPWork&,QPCtx& c,real curFarBound){ c.farbound=curFarBound; c.farRampOffset=c.rampOffset; }
// QProblem::solveQP (QProblem.cpp:1477-1747); nWSR in/out like the reference
// PARK (wbc_solve_cold_kernel only): the gang's barrier also tells how many of its threads are done; as soon as a member has finished -- and the cheap early iterations
// are over -- the solve returns QP_SUSPENDED from the top of the iteration, to be continued from exactly there (c.resuming) by another launch.
template<bool PARK> WBC_FN int solveQP(QPWork& w,QPCtx& c,int& nWSR,int nWSRperformed){
    if(PARK&&c.resuming){ c.resuming=0; } else {
    if(c.status==QPS_NOTINITIALISED||c.status==QPS_PREPARINGAUXILIARYQP||c.status==QPS_PERFORMINGHOMOTOPY) return QP_RET_HOTSTART_FAILED_AS_QP_NOT_INITIALISED;
    c.status=QPS_PERFORMINGHOMOTOPY;                 // updateActivitiesForHotstart: nothing to do, every far bound is finite
    c.infeasible=0; c.unbounded=0;
    }
    for(int iter=nWSRperformed;iter<nWSR;iter++){
        c.status=QPS_PERFORMINGHOMOTOPY;
#if !defined(__CUDACC__) && defined(WBC_QP_DBG)
        g_dbg_iter=iter;
#endif
#if defined(WBC_GANG) && defined(__CUDACC__)
        if(PARK){ if(gang_count(0)>=32*PARK_MIN_DONE&&c.parkOK&&iter>=PARK_MIN_ITER){ c.park_iter=iter; nWSR=iter; return QP_SUSPENDED; } }
        else
#endif
        WBC_GANG_SYNC();
        WBC_TIC(t0);
        bool bCzero,bBzero; determineDataShift(w,c,bCzero,bBzero);
        WBC_TOC(t0,0);
        int rv=determineStepDirection(w,c,nullptr,bCzero,bBzero);
        WBC_TOC(t0,1);
        if(rv!=QP_OK){ nWSR=iter; return rv; }
        int BC_idx,BC_status; bool BC_isBound;
        performStep(w,c,BC_idx,BC_status,BC_isBound);
        const bool finished=homotopyFinished(w,c);
        WBC_TOC(t0,2); WBC_COUNT_ITER();
        if(finished){ c.status=QPS_SOLVED; nWSR=iter; return QP_OK; }
#if !defined(__CUDACC__) && defined(WBC_QP_TRACE)
        printf("  it %3d tau %.6e  %s %s %3d  nFX %2d nAC %2d\n",iter,c.tau,BC_status==2?"---":(BC_status==0?"REM":(BC_status<0?"ADL":"ADU")),BC_isBound?"B":"C",BC_idx,c.nFX,c.nAC);
#endif
        rv=changeActiveSet(w,c,BC_idx,BC_status,BC_isBound);
        WBC_TOC(t0,3);
        if(rv!=QP_OK){ nWSR=iter; if(c.infeasible){ c.status=QPS_HOMOTOPYQPSOLVED; return QP_RET_HOTSTART_STOPPED_INFEASIBILITY; } return rv; }
        rv=(BC_status==ST_INACTIVE&&c.nAC>0&&c.nFR>c.nAC)? extendProjectedCholesky(w,c) : computeProjectedCholesky(w,c);
        WBC_TOC(t0,4);
        if(rv!=QP_OK) return rv;
        c.status=QPS_HOMOTOPYQPSOLVED;
        if(BC_status!=2&&c.tau<=QEPS) performRamping(w,c); else performDriftCorrection(w,c);
        WBC_TOC(t0,5);
    }
    return QP_RET_MAX_NWSR_REACHED;
}
// QProblem::hotstart (QProblem.cpp:446-640): far-bound outer loop
template<bool PARK> WBC_FN int hotstartFar(QPWork& w,QPCtx& c,int& nWSR){
    const int lane=WBC_LANE(c); int nWSR_max=nWSR; int nWSR_performed=0;
    real farbound=OPT_initialFarBounds;
    if(PARK&&c.resuming){ nWSR_max=c.park_nWSRmax; nWSR_performed=c.park_iter; farbound=c.park_farbound; }      // c.farbound / c.farRampOffset came back with the context
    else {
    { real m=real(0.0); LANE_FOR(i,70){ real u=(real)c.qg->ubA[i], l=(real)c.qg->lbA[i]; if(u<QINFTY) m=fmax(m,u); if(l>-QINFTY) m=fmax(m,-l); } m=wbc_max(m); if(m>farbound) farbound=m; }
    updateFarBounds(w,c,farbound);
    }
    int rv;
    for(;;){
        nWSR=nWSR_max;
        rv=solveQP<PARK>(w,c,nWSR,nWSR_performed);
        if(PARK&&rv==QP_SUSPENDED){ c.park_nWSRmax=nWSR_max; c.park_farbound=farbound; return rv; }
        nWSR_performed=nWSR;
        farbound*=OPT_growFarBounds;
        if(c.infeasible){
            if(farbound>=QINFTY){ rv=QP_RET_HOTSTART_STOPPED_INFEASIBILITY; break; }
            updateFarBounds(w,c,farbound);
        } else if(c.status==QPS_SOLVED){
            const real tol=farbound/OPT_growFarBounds*OPT_boundTolerance; int na=0;
            LANE_FOR(i,42){ if(fabs(tgt_lb(c,i)-w.x[i])<tol) na++; if(fabs(tgt_ub(c,i)-w.x[i])<tol) na++; }
            LV_FOR(i,sl,70){ LV_LD(axi,Ax,i,sl); LV_WAIT(); LV_USE(axi);
                if(LV_IN(i,70)){ const real l=tgt_lbA(w,lv_h_,i), u=tgt_ubA(w,lv_h_,i); if(l>(real)lv_h_.qg->lbA[i]&&fabs(l-axi)<tol) na++; if(u<(real)lv_h_.qg->ubA[i]&&fabs(u-axi)<tol) na++; } }
            if(wbc_ballot(na>0)==0u) break;
            c.status=QPS_HOMOTOPYQPSOLVED;
            if(farbound>=QINFTY){ c.unbounded=1; rv=QP_RET_HOTSTART_STOPPED_UNBOUNDEDNESS; break; }
            updateFarBounds(w,c,farbound);
        } else break;
        c.rampOffset++;
    }
    return rv;
}
// SQProblem/QProblem::init -> solveInitialQP (QProblem.cpp:316-368, 1301-1471): cold start from x = y = 0 with every variable
// fixed at its (far) lower bound (initialStatusBounds = ST_LOWER) and no active constraint.
template<bool PARK> WBC_FN int qp_init(QPWork& w,QPCtx& c,int& nWSR){
    const int lane=WBC_LANE(c); WBC_SYNC();
    int rv;
    if(!(PARK&&c.resuming)){
    LANE_FOR(i,42){ w.x[i]=real(0.0); w.y[i]=real(0.0); w.FX[i]=(signed char)i; w.bstat[i]=ST_LOWER; }
    LV_FOR(i,sl,42){ LV_ST(g,i,sl,real(0.0)); LV_ST(lb,i,sl,real(0.0)); LV_ST(ub,i,sl,OPT_boundRelaxation); LV_DONE(); }
    LANE_FOR(i,70){ w.y[42+i]=real(0.0); w.IAC[i]=(signed char)i; w.cstat[i]=ST_INACTIVE; w.posAC[i]=0; }
    LV_FOR(i,sl,70){ LV_ST(Ax,i,sl,real(0.0)); LV_ST(lbA,i,sl,-OPT_boundRelaxation); LV_ST(ubA,i,sl,OPT_boundRelaxation); LV_ST(Ax_l,i,sl,OPT_boundRelaxation); LV_ST(Ax_u,i,sl,OPT_boundRelaxation); LV_DONE(); }
    c.nFR=0; c.nFX=42; c.nAC=0; c.nIAC=70; c.frIdentity=0; c.rampOffset=0; c.infeasible=0; c.unbounded=0; c.tau=real(0.0);
    setupTQfactorisation(w,c);
    c.status=QPS_AUXILIARYQPSOLVED;
    performRamping(w,c);
    rv=computeProjectedCholesky(w,c);                        // setupInitialCholesky
    if(rv!=QP_OK) return QP_RET_INIT_FAILED_CHOLESKY;
    }
    rv=hotstartFar<PARK>(w,c,nWSR);
    if(PARK&&rv==QP_SUSPENDED) return rv;
    if(c.infeasible) return QP_RET_INIT_FAILED_INFEASIBILITY;
    if(c.unbounded) return QP_RET_INIT_FAILED_UNBOUNDEDNESS;
    if(rv!=QP_OK&&rv!=QP_RET_MAX_NWSR_REACHED) return QP_RET_INIT_FAILED_HOTSTART;
    return rv;
}
// The TQ part of SQProblem::setupNewAuxiliaryQP (SQProblem.cpp:340-515 -> setupAuxiliaryWorkingSet, QProblem.cpp:2212-2378): everything
// free / inactive again, Q = I, T empty, then the old active constraints (`old` = their statuses) re-added in ascending order with the
// cheap LI test and no LI exchange.  Touches only the prefix of the workspace (Q, T, lists, Givens vectors, arow) and reads the QP
// data through `qd`, so it also runs as a kernel of its own with 9 robots per SM (wbc_warm_setup_kernel).
template<class QD> WBC_FN void rebuild_TQ(QPWork& w,QPCtx& c,const QD& qd,const signed char* old){
    const int lane=WBC_LANE(c);
    LANE_FOR(i,42){ w.FR[i]=(signed char)i; }
    LANE_FOR(i,70){ w.IAC[i]=(signed char)i; w.cstat[i]=ST_INACTIVE; }
    WBC_SYNC();
    c.nFR=42; c.nFX=0; c.nAC=0; c.nIAC=70; c.frIdentity=1;
    setupTQfactorisation(w,c);
    for(int i=0;i<70;i++){
        const int os=(int)old[i];
        if(os==ST_INACTIVE) continue;
        setup_readd_constraint(w,c,qd,i,os);
    }
}
// store / load of the rebuilt factorisation (HBM, coalesced: consecutive lanes, consecutive doubles).  Loads are issued in batches ahead of the
// shared stores (see the prologue of wbc_solve_qp_kernel).
WBC_DEV void store_prebuilt_TQ(const QPWork& w,const QPCtx& c,QPPre& p){
    const int lane=WBC_LANE(c), nT=tri(c.nAC);
    LANE_FOR(i,42*LD) p.Q[i]=w.Q[i];
    LANE_FOR(i,nT) p.T[i]=w.RT[i];
    LANE_FOR(i,70){ p.AC[i]=w.AC[i]; p.IAC[i]=w.IAC[i]; p.cstat[i]=w.cstat[i]; p.posAC[i]=w.posAC[i]; }
    if(lane==0){ p.nAC=c.nAC; p.nIAC=c.nIAC;
#if defined(WBC_PROFILE)
        p.flops=c.flops;
#endif
    }
}
WBC_DEV void load_prebuilt_TQ(QPWork& w,QPCtx& c,const QPPre& p){
    const int lane=WBC_LANE(c);
    c.nFR=42; c.nFX=0; c.nAC=p.nAC; c.nIAC=p.nIAC; c.frIdentity=1; WBC_FLOPS(p.flops);
    const int nT=tri(c.nAC);
#if WBC_NL==1
    LANE_FOR(i,70){ w.AC[i]=p.AC[i]; w.IAC[i]=p.IAC[i]; w.cstat[i]=p.cstat[i]; w.posAC[i]=p.posAC[i]; }
#else
    signed char l0[3],l1[3],l2[3],l3[3];
    #pragma unroll
    for(int k=0;k<3;k++){ const int i=lane+WBC_NL*k; const bool in=i<70; l0[k]=in?p.AC[i]:(signed char)0; l1[k]=in?p.IAC[i]:(signed char)0; l2[k]=in?p.cstat[i]:(signed char)0; l3[k]=in?p.posAC[i]:(signed char)0; }
#endif
#if WBC_NL>1
    if(c.tqBar){                                                               // Q and T were requested by a bulk copy when the robot was fetched (wbc_solve_qp_kernel): wait for it
        wbc_mbar_wait(c.tqBar,c.tqParity); c.tqBar=0u;
        constexpr int QT=42*LD+TRI42+1, QT16=(int)((QT*sizeof(real))/16*16/sizeof(real));     // fp32: the last two elements do not fill a 16-byte unit
        if(QT16<QT&&lane<QT-QT16&&QT16+lane-42*LD<nT) w.RT[QT16+lane-42*LD]=p.T[QT16+lane-42*LD];
    } else
#endif
    {
    constexpr int KB=19;                                                        // 3 batches of 19 x 32 doubles cover Q (1806)
    #pragma unroll 1
    for(int base=0;base<42*LD;base+=KB*WBC_NL){
        real t[KB];
        #pragma unroll
        for(int k=0;k<KB;k++){ const int i=base+lane+WBC_NL*k; t[k]=(i<42*LD)?p.Q[i]:real(0.0); }
        #pragma unroll
        for(int k=0;k<KB;k++){ const int i=base+lane+WBC_NL*k; if(i<42*LD) w.Q[i]=t[k]; }
    }
    #pragma unroll 1
    for(int base=0;base<nT;base+=KB*WBC_NL){
        real t[KB];
        #pragma unroll
        for(int k=0;k<KB;k++){ const int i=base+lane+WBC_NL*k; t[k]=(i<nT)?p.T[i]:real(0.0); }
        #pragma unroll
        for(int k=0;k<KB;k++){ const int i=base+lane+WBC_NL*k; if(i<nT) w.RT[i]=t[k]; }
    }
    }
#if WBC_NL>1
    #pragma unroll
    for(int k=0;k<3;k++){ const int i=lane+WBC_NL*k; if(i<70){ w.AC[i]=l0[k]; w.IAC[i]=l1[k]; w.cstat[i]=l2[k]; w.posAC[i]=l3[k]; } }
#endif
    LANE_FOR(i,42) w.FR[i]=(signed char)i;
    WBC_SYNC();
}
// SQProblem::hotstart -> setupNewAuxiliaryQP (SQProblem.cpp:153-197, 340-515) then QProblem::hotstart.
// On entry w.x, w.y, w.lb, w.ub, w.Ax_l, w.Ax_u, w.bstat, w.cstat hold the previous tick's solver state.
WBC_FN int qp_hotstart(QPWork& w,QPCtx& c,int& nWSR,const QPPre* pre){
    const int lane=WBC_LANE(c);
    if(c.status==QPS_NOTINITIALISED||c.status==QPS_PREPARINGAUXILIARYQP||c.status==QPS_PERFORMINGHOMOTOPY) return QP_RET_HOTSTART_FAILED_AS_QP_NOT_INITIALISED;
    c.status=QPS_PREPARINGAUXILIARYQP; WBC_SYNC();
    WBC_FLOPS(4.0*42*42+4.0*70*42);
    LV_FOR(i,sl,70){ LV_LD(al,Ax_l,i,sl); LV_LD(au,Ax_u,i,sl); LV_WAIT(); LV_USE(al); LV_USE(au);
        real ax=real(0.0), l=real(0.0), u=real(0.0);
        if(LV_IN(i,70)){ ax=Arow_acc(WQP,i,w.x,real(0.0),real(1.0)); l=-al+ax; u=au+ax; au=u-ax; al=ax-l; }
        LV_ST(Ax,i,sl,ax); LV_ST(lbA,i,sl,l); LV_ST(ubA,i,sl,u); LV_ST(Ax_u,i,sl,au); LV_ST(Ax_l,i,sl,al); LV_DONE(); }
    setupAuxiliaryQPgradient(w,c);
    // old working set (statuses), then everything free / inactive again
    bool anyFixed=false; LANE_FOR(i,42) if(w.bstat[i]!=0) anyFixed=true;
    if(wbc_ballot(anyFixed)!=0u) return QP_RET_SETUP_AUXILIARYQP_FAILED;     // addBound on an ST_UNBOUNDED-typed variable: RET_ALL_BOUNDS_ACTIVE
    if(!pre){
        signed char* old=reinterpret_cast<signed char*>(w.tmpv);                 // park the old statuses (nothing in the rebuild touches tmpv)
        LANE_FOR(i,70) old[i]=w.cstat[i];
        WBC_SYNC();
        rebuild_TQ(w,c,WQP,old);
    } else load_prebuilt_TQ(w,c,*pre);                                         // the same, done beforehand by wbc_warm_setup_kernel
    int rv=computeProjectedCholesky(w,c);
    if(rv!=QP_OK) return QP_RET_UNSUPPORTED_PATH;
    c.status=QPS_AUXILIARYQPSOLVED;
    return hotstartFar<false>(w,c,nWSR);
}

}  // namespace wbc
