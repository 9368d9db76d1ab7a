// DEVELOPMENT AID ONLY -- not part of the shipped library (see emul_rbd.cpp).  One emulated lane.
#include "../../quadrupedwholebodycontroller_b200/csrc/wbc_qp.cuh"
#include <cstring>
#include <cstdlib>
#include <cstdio>
using namespace wbc;

struct EmulQP { QPWork w; QPCtx c; QPData qg; };
extern "C" void* emulqp_create(){ EmulQP* e=new EmulQP; memset(e,0,sizeof(*e)); e->c.lane=0; e->c.status=QPS_NOTINITIALISED; return e; }
extern "C" void emulqp_destroy(void* h){ delete (EmulQP*)h; }
// first != 0: init (cold), else hotstart.  outputs: x[42], y[112], ws[112], returns rc; *nwsr in/out
extern "C" int emulqp_solve(void* h,const double* qpdata,int first,int* nwsr,double* x,double* y,signed char* ws){
    EmulQP* e=(EmulQP*)h; { real* dst=(real*)&e->w.qp; for(int i=0;i<QPHOT_HEAD;i++) dst[i]=(real)qpdata[i]; for(int i=0;i<6;i++) dst[QPHOT_HEAD+i]=(real)qpdata[QPDATA_DOUBLES-6+i];
      memcpy(&e->qg,qpdata,sizeof(QPData)); e->c.qg=&e->qg; }
    int n=*nwsr; int rc;
    if(first) rc=qp_init<false>(e->w,e->c,n);
    else if(getenv("EMUL_SPLIT")){          // the two-kernel warm start: rebuild in a second workspace, hand over through QPPre
        static EmulQP a; static QPPre pre; memset(&a,0,sizeof(a)); memset(&pre,0,sizeof(pre)); a.c.lane=0; a.c.qg=&e->qg;
        signed char old[70]; memcpy(old,e->w.cstat,70);
        rebuild_TQ(a.w,a.c,e->w.qp,old); store_prebuilt_TQ(a.w,a.c,pre);
        rc=qp_hotstart(e->w,e->c,n,&pre);
    } else rc=qp_hotstart(e->w,e->c,n,nullptr);
    *nwsr=n;
    for(int i=0;i<42;i++) x[i]=e->w.x[i]; for(int i=0;i<112;i++) y[i]=e->w.y[i];
    for(int i=0;i<42;i++) ws[i]=e->w.bstat[i]; for(int i=0;i<70;i++) ws[42+i]=e->w.cstat[i];
    return rc;
}
