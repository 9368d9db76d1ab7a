// DEVELOPMENT AID ONLY -- not part of the shipped library, never loaded by the product path.
// Compiles the CUDA device source (csrc/*.cuh) as plain C++ with ONE emulated lane (WBC_NL == 1) so that control
// flow and algebra can be debugged against the oracle on a machine without a GPU.
#include "../../quadrupedwholebodycontroller_b200/csrc/wbc_assembly.cuh"
#include "../../quadrupedwholebodycontroller_b200/csrc/anymal_model.h"
#include <cstring>
using namespace wbc;

static void fill_model(ModelConst& m){
    memcpy(m.restR,ANYMAL_restR,sizeof(m.restR)); memcpy(m.restP,ANYMAL_restP,sizeof(m.restP)); memcpy(m.axis,ANYMAL_axis,sizeof(m.axis));
    memcpy(m.mass,ANYMAL_mass,sizeof(m.mass)); memcpy(m.mcom,ANYMAL_mcom,sizeof(m.mcom)); memcpy(m.inertia,ANYMAL_inertia,sizeof(m.inertia));
    memcpy(m.footR,ANYMAL_footR,sizeof(m.footR)); memcpy(m.footP,ANYMAL_footP,sizeof(m.footP));
    m.total_mass=0; for(int i=0;i<13;i++) m.total_mass+=m.mass[i];
}
extern "C" void emul_tick(const double* params16,const double* q,const double* qd,const double* pos,const double* quat,const double* twist,
                          const double* pose,const double* vel,const double* acc,const double* swp,const double* swv,const double* swa,
                          const unsigned char* contacts,int switching,int first,double* state /*CTRLSTATE_DOUBLES*/,double* qpdata /*QPDATA_DOUBLES*/){
    static ModelConst mc; static bool init=false; if(!init){ fill_model(mc); init=true; }
    Params P; memcpy(&P,params16,sizeof(P));
    static RbdScratch s;
    for(int i=0;i<12;i++){ s.q[i]=q[i]; s.qd[i]=qd[i]; s.swp[i]=swp[i]; s.swv[i]=swv[i]; s.swa[i]=swa[i]; }
    for(int i=0;i<6;i++){ s.twist[i]=twist[i]; s.pose_d[i]=pose[i]; s.vel_d[i]=vel[i]; s.acc_d[i]=acc[i]; }
    for(int i=0;i<3;i++) s.pb[i]=pos[i];
    for(int i=0;i<4;i++) s.cmask[i]=contacts[i]?1.0:0.0;
    quat_to_R(quat,s.Rwb);
    mtv3(s.Rwb,s.twist,s.vbody); mtv3(s.Rwb,s.twist+3,s.vbody+3);
    for(int leg=0;leg<4;leg++) rbd_leg(mc,s,leg,leg);
    wbc_assemble(mc,P,s,*(CtrlState*)state,*(QPData*)qpdata,*(QPData*)qpdata,0,switching,first,nullptr,1);
}
extern "C" int emul_sizes(int which){ return which==0?CTRLSTATE_DOUBLES:QPDATA_DOUBLES; }
