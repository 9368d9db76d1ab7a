// Shared definitions for the batched ANYmal whole-body-controller kernels (sm_100a).
//
// Execution model: ONE ROBOT INSTANCE PER WARP.  Per-instance matrices live in shared memory, scalars that steer
// control flow are replicated in registers of all 32 lanes (every lane takes the same branches), vector work is
// split as "lane i owns element i, i+32, i+64" and inner products run sequentially inside a lane so that sums are
// accumulated in the same order as the reference's scalar loops.
//
// The same source also compiles as plain C++ with WBC_NL == 1 (one "lane") -- used ONLY by tools/host_emul to
// debug control flow against the oracle on a machine without a GPU; it is never part of the shipped library.
#pragma once
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
  #define WBC_DEV __device__ __forceinline__
  #define WBC_HD  __host__ __device__ __forceinline__
  #define WBC_FN  __device__ __noinline__
  #define WBC_NL 32
  #define WBC_SYNC() __syncwarp()
  // explicit shared-space loads with a 32-bit address held in a register: through C++ pointers ptxas re-derives the shared window
  // base (S2R SR_CgaCtaId + MOV + LEA) at every predicated access inside the noinline routines.  READ-ONLY data of the calling routine only
  // (the asm is not volatile and has no memory clobber, so the compiler may move or merge these loads).
  WBC_DEV unsigned wbc_saddr(const void* p){ unsigned a=(unsigned)__cvta_generic_to_shared(p); asm volatile("" : "+r"(a)); return a; }   // opaque: stays in a register
#if defined(WBC_FP32)
  WBC_DEV float wbc_lds(unsigned a){ float v; asm("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a)); return v; }
  WBC_DEV float wbc_max(float v){ for(int o=16;o>0;o>>=1){ float t=__shfl_xor_sync(0xffffffffu,v,o); v = t>v?t:v; } return v; }
  WBC_DEV float wbc_bcast(float v,int src){ return __shfl_sync(0xffffffffu,v,src); }
#else
  WBC_DEV double wbc_lds(unsigned a){ double v; asm("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a)); return v; }
#endif
  // Bulk asynchronous copies HBM -> shared memory (cp.async.bulk, SASS UBLKCP) completing on an mbarrier: one instruction moves a robot's
  // rebuilt factorisation (Q and T, up to 21.7 KB) or its QP blocks (4.4 KB) while the warp goes on with other work, instead of staging
  // 32-element rounds through registers with a memory round trip each.  Addresses and sizes are multiples of 16 bytes.
  WBC_DEV void wbc_mbar_init(unsigned bar,unsigned count){ asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(bar),"r"(count) : "memory"); }
  WBC_DEV void wbc_mbar_init_fence(){ asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  WBC_DEV void wbc_fence_proxy_async(){ asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }      // generic-proxy accesses of this thread before async-proxy writes
  WBC_DEV void wbc_mbar_expect_tx(unsigned bar,unsigned bytes){ asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar),"r"(bytes) : "memory"); }
  WBC_DEV void wbc_bulk_g2s(unsigned dst,const void* src,unsigned bytes,unsigned bar){
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" :: "r"(dst),"l"(__cvta_generic_to_global(src)),"r"(bytes),"r"(bar) : "memory"); }
  WBC_DEV void wbc_mbar_wait(unsigned bar,unsigned parity){
      unsigned ok; do{ asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(bar),"r"(parity) : "memory"); }while(!ok); }
  // Tensor memory as a LANE-PRIVATE scratchpad (WBC_TMEM): thread t of warp w owns TMEM lane 32*(w%4)+t; one `real` is one (fp32) or two (fp64) 32-bit
  // columns, moved by tcgen05.ld / tcgen05.st 32x32b (SASS LDTM / STTM).  A load's registers are valid after tcgen05.wait::ld (wbc_tm_wait_ld, then
  // wbc_tm_use on each value so the compiler keeps every use behind the wait); stores are complete after tcgen05.wait::st.  All of them are .sync.aligned:
  // every lane of the warp executes them together.
#if defined(WBC_FP32)
  WBC_DEV void wbc_tm_ld(unsigned a,float& v){ unsigned x; asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(x) : "r"(a)); v=__uint_as_float(x); }
  WBC_DEV void wbc_tm_st(unsigned a,float v){ asm volatile("tcgen05.st.sync.aligned.32x32b.x1.b32 [%0], {%1};" :: "r"(a),"r"(__float_as_uint(v)) : "memory"); }
  WBC_DEV void wbc_tm_use(float& v){ asm volatile("" : "+f"(v)); }
  constexpr unsigned WBC_TM_COLS_PER_REAL=1u;
#else
  WBC_DEV void wbc_tm_ld(unsigned a,double& v){ unsigned lo,hi; asm volatile("tcgen05.ld.sync.aligned.32x32b.x2.b32 {%0, %1}, [%2];" : "=r"(lo),"=r"(hi) : "r"(a)); v=__hiloint2double((int)hi,(int)lo); }
  WBC_DEV void wbc_tm_st(unsigned a,double v){ asm volatile("tcgen05.st.sync.aligned.32x32b.x2.b32 [%0], {%1, %2};" :: "r"(a),"r"((unsigned)__double2loint(v)),"r"((unsigned)__double2hiint(v)) : "memory"); }
  WBC_DEV void wbc_tm_use(double& v){ asm volatile("" : "+d"(v)); }
  constexpr unsigned WBC_TM_COLS_PER_REAL=2u;
#endif
  WBC_DEV void wbc_tm_wait_ld(){ asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
  WBC_DEV void wbc_tm_wait_st(){ asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
  WBC_DEV unsigned wbc_ballot(bool p){ return __ballot_sync(0xffffffffu,p); }
  WBC_DEV double wbc_max(double v){ for(int o=16;o>0;o>>=1){ double t=__shfl_xor_sync(0xffffffffu,v,o); v = t>v?t:v; } return v; }
  WBC_DEV double wbc_bcast(double v,int src){ return __shfl_sync(0xffffffffu,v,src); }
  WBC_DEV int    wbc_bcasti(int v,int src){ return __shfl_sync(0xffffffffu,v,src); }
  WBC_DEV int    wbc_ffs(unsigned m){ return __ffs(m)-1; }
  WBC_DEV int    wbc_popc(unsigned m){ return __popc(m); }
  WBC_DEV int    wbc_rank(unsigned m,int lane){ return __popc(m&((1u<<lane)-1u)); }   // set bits below `lane`
#else
  #define WBC_DEV static inline
  #define WBC_HD  static inline
  #define WBC_FN  static
  #define WBC_NL 1
  #define WBC_SYNC() do{}while(0)
  WBC_DEV unsigned wbc_ballot(bool p){ return p?1u:0u; }
  WBC_DEV double wbc_max(double v){ return v; }
  WBC_DEV double wbc_bcast(double v,int){ return v; }
  WBC_DEV float wbc_max(float v){ return v; }
  WBC_DEV float wbc_bcast(float v,int){ return v; }
  WBC_DEV int    wbc_bcasti(int v,int){ return v; }
  WBC_DEV int    wbc_ffs(unsigned m){ return m?0:-1; }
  WBC_DEV int    wbc_popc(unsigned m){ return m?1:0; }
  WBC_DEV int    wbc_rank(unsigned,int){ return 0; }
#endif

// element loops are NOT unrolled: the solver is instruction-cache bound (453 KB of code vs a 32 KB L1.5), not issue bound
#if defined(__CUDACC__) && defined(WBC_LANE_UNROLL)
  #define WBC_PRAGMA_(x) _Pragma(#x)
  #define WBC_PRAGMA(x) WBC_PRAGMA_(x)
  #define LANE_FOR(i,n) WBC_PRAGMA(unroll WBC_LANE_UNROLL) for(int i=lane;i<(n);i+=WBC_NL)
#elif defined(__CUDACC__)
  #define LANE_FOR(i,n) _Pragma("unroll 1") for(int i=lane;i<(n);i+=WBC_NL)
#else
  #define LANE_FOR(i,n) for(int i=lane;i<(n);i+=WBC_NL)
#endif

namespace wbc {

// arithmetic type of the QP solver: fp64 like qpOASES' default build, or fp32 (-DWBC_FP32 -> libwbc_b200_f32.so, the north star's optional mode:
// qpOASES' __USE_SINGLE_PRECISION__ constants and tolerances).  RBD, assembly, planner and plant always run in fp64.
#if defined(WBC_FP32)
  typedef float real;
#else
  typedef double real;
#endif

constexpr int NJ=12, NLEG=4, NV=42, NC=70;              // whole_body_controller.hpp:27-32
constexpr double GRAV=9.81;

// gains and limits shared by the whole batch (whole_body_controller.hpp:91-104, config/params_controller.yaml)
struct Params {
    double friction, loopRate, maxTorque, kpValue, kpValueZ, kdValue, kiValue, kpSwingValue, kdSwingValue, slackWeight;
    double refPose[6];
};

// ---- compact QP handed from the assembly stage to the solver (doubles, per instance) ----
// The dense 42x42 H and 70x42 A of whole_body_controller.cpp:479-497 have few distinct blocks:
//   H = blkdiag(I6, I12, Hff, rho*I12),  Hff = I12 + Jst_b Jst_b^T
//   A rows 0:6   [Mbb 0 -Jst_b^T 0]      rows 6:18  [Jst_b Jst_j 0 0]     rows 18:34 [0 0 Dfr 0]
//     rows 34:46 [0 Mjj -Jst_j^T 0]      rows 46:58 [Jsw_b Jsw_j 0 -I]    rows 58:70 [Jsw_b Jsw_j 0 +I]
// with Jst = c (.) Jc and Jsw = (1-c) (.) Jc row masks of ONE centroidal foot Jacobian Jc (12x18).
template<class T> struct QPDataT {
    typedef T value_type;
    T Jc[12*18];      // centroidal foot Jacobian (all four feet, unmasked), row-major
    T Mbb[36];        // centroidal mass matrix, base block
    T Mjj[144];       // centroidal mass matrix, joint block
    T Hff[144];       // H[18:30,18:30]
    T gf[12];         // g[18:30]
    T lbA[70], ubA[70];
    T biasj[12];      // centroidal bias, joint part (for the torque recovery, wbc.cpp:565-568)
    T mu, rho;
    T cmask[4];       // 1.0 = stance, 0.0 = swing  (leg order LH, LF, RF, RH)
};
typedef QPDataT<double> QPData;      // what the assembly kernel writes to HBM (always fp64)
// The part of it the solver keeps in shared memory, in its arithmetic type: the matrix blocks every product with A and H reads.  The constraint
// bounds lbA / ubA (read twice per working-set change, element-wise) and bias_j (read once, by the torque epilogue) stay in HBM / L1: 1.2 KB of
// shared memory per robot that decide whether an SM holds 5 or 6 robots.
template<class T> struct QPHotT {
    typedef T value_type;
    T Jc[12*18], Mbb[36], Mjj[144], Hff[144], gf[12];    // same order as the head of QPDataT
    T mu, rho, cmask[4];
};
typedef QPHotT<real> QPHotR;
constexpr int QPHOT_HEAD = 12*18+36+144+144+12;            // leading elements shared with QPDataT
static_assert(sizeof(QPHotT<double>)/sizeof(double)==QPHOT_HEAD+6,"QPHot layout");
constexpr int QPDATA_DOUBLES = sizeof(QPData)/sizeof(double);

// ---- per-instance persistent controller state (wbc.hpp:150-161) kept in HBM between ticks ----
struct CtrlState {
    double Told[6*18];     // top 6 rows of oldTransformationMatrix_ (rows 6..17 are constant [0 I])
    double Jstold[12*18];  // oldCentroidStanceJacobian_
    double Jswold[12*18];  // oldCentroidSwingJacobian_
    double TdotInv[6*18];  // top 6 rows of transformationMatrixDotInverse_ (rows 6..17 are identically zero)
    double integral[6];    // integralError_
};
constexpr int CTRLSTATE_DOUBLES = sizeof(CtrlState)/sizeof(double);

// ---- small fixed-size helpers (3-vectors, 3x3 row-major) ----
WBC_DEV void cross3(const double* a,const double* b,double* r){ double x=a[1]*b[2]-a[2]*b[1], y=a[2]*b[0]-a[0]*b[2], z=a[0]*b[1]-a[1]*b[0]; r[0]=x;r[1]=y;r[2]=z; }
WBC_DEV void mv3(const double* R,const double* v,double* r){ double x=R[0]*v[0]+R[1]*v[1]+R[2]*v[2], y=R[3]*v[0]+R[4]*v[1]+R[5]*v[2], z=R[6]*v[0]+R[7]*v[1]+R[8]*v[2]; r[0]=x;r[1]=y;r[2]=z; }
WBC_DEV void mtv3(const double* R,const double* v,double* r){ double x=R[0]*v[0]+R[3]*v[1]+R[6]*v[2], y=R[1]*v[0]+R[4]*v[1]+R[7]*v[2], z=R[2]*v[0]+R[5]*v[1]+R[8]*v[2]; r[0]=x;r[1]=y;r[2]=z; }
WBC_DEV void mm3(const double* A,const double* B,double* C){ double t[9]; for(int i=0;i<3;i++)for(int j=0;j<3;j++) t[3*i+j]=A[3*i]*B[j]+A[3*i+1]*B[3+j]+A[3*i+2]*B[6+j]; for(int i=0;i<9;i++) C[i]=t[i]; }
WBC_DEV double dot3(const double* a,const double* b){ return a[0]*b[0]+a[1]*b[1]+a[2]*b[2]; }

}  // namespace wbc
