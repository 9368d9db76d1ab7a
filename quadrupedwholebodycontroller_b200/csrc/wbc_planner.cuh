// Batched MotionPlanner: the producer of the controller's references, one planner per robot.
//
// Replaces the reference's MotionPlanner (src/motion_planner.cpp:130-177 constructor, :180-376 plannerLoop, :5-97 Bezier and
// quintic helpers; config/params_planner.yaml) and the hand-over of its message to the controller
// (WholeBodyController::referenceCallback, src/whole_body_controller.cpp:150-185).
// The reference's planner is a blocking nested loop paced by ros::Rate; here it is a resumable state machine advanced by one
// planner period per call (one thread per robot), so that a closed-loop rollout never leaves the GPU:
//   mode IDLE       top of the outer loop: a non-zero /cmd_vel starts a cycle (foot targets advance), else "stand" is published
//   mode IN_CYCLE   one pass of the inner loop: publish a sample of the running step phase, or (no message) move to the next phase
//   mode CYCLE_END  the period after the 4th phase ended: cycle bookkeeping, no message
// A published message is latched into the controller's reference buffers and isSwitchingFootState_ is re-derived from the
// previous message's contacts exactly like referenceCallback; a period without message leaves both untouched.
// Products and sums use __dmul_rn / __dadd_rn so that no FMA contraction happens: results are bit-identical to the scalar
// CPU evaluation order of the reference (only sin/cos of a non-zero yaw may differ in the last place).
#pragma once
#include "wbc_common.cuh"

namespace wbc {

#if defined(__CUDACC__)
  WBC_DEV double pmul(double a,double b){ return __dmul_rn(a,b); }
  WBC_DEV double padd(double a,double b){ return __dadd_rn(a,b); }
#else
  WBC_DEV double pmul(double a,double b){ return a*b; }
  WBC_DEV double padd(double a,double b){ return a+b; }
#endif
WBC_DEV double psub(double a,double b){ return padd(a,-b); }

// config/params_planner.yaml + the three quintic time laws of plannerLoop (motion_planner.cpp:190-195), evaluated once on the host
struct PlannerParams {
    double step_length, height_control_point, x_offset, y_offset, step_duration, cycle_duration, body_height, body_final_velocity, dt;
    double poly_foot[6], poly_body_start[6], poly_body_continue[6];
};
enum : int { PL_IDLE=0, PL_IN_CYCLE=1, PL_CYCLE_END=2 };

// per-robot planner state (motion_planner.hpp:22-38 members + the locals of plannerLoop + the latched ref_msg_)
struct PlannerState {
    double cmd[3];                      // velocity_command_(0), velocity_command_(1), yaw_rate_command_
    double yaw, step_time, cycle_time;  // yaw_, current_step_time, current_cycle_time
    double vrot[3];                     // velocity_command_rotated of the running cycle
    double pi_body[3], pf_body[3];
    double pi_foot[4][3], pf_foot[4][3];        // leg order of the message: LH, LF, RF, RH
    double pose[6], vel[6], acc[6], swp[12], swv[12], swa[12];   // ref_msg_
    int step_phase, cycle_counter, mode, pad_;
    unsigned char contacts[4], pad2_[4];
};

// quintic time law (motion_planner.cpp:59-75), left-to-right like the reference's expressions
WBC_DEV double quintic(const double* a,double t){
    double r=padd(a[0],pmul(a[1],t)); r=padd(r,pmul(pmul(a[2],t),t)); r=padd(r,pmul(pmul(pmul(a[3],t),t),t));
    r=padd(r,pmul(pmul(pmul(pmul(a[4],t),t),t),t)); return padd(r,pmul(pmul(pmul(pmul(pmul(a[5],t),t),t),t),t));
}
WBC_DEV double quintic_d1(const double* a,double t){
    double r=padd(a[1],pmul(pmul(2.0,a[2]),t)); r=padd(r,pmul(pmul(pmul(3.0,a[3]),t),t)); r=padd(r,pmul(pmul(pmul(pmul(4.0,a[4]),t),t),t));
    return padd(r,pmul(pmul(pmul(pmul(pmul(5.0,a[5]),t),t),t),t));
}
WBC_DEV double quintic_d2(const double* a,double t){
    double r=padd(pmul(2.0,a[2]),pmul(pmul(6.0,a[3]),t)); r=padd(r,pmul(pmul(pmul(12.0,a[4]),t),t)); return padd(r,pmul(pmul(pmul(pmul(20.0,a[5]),t),t),t));
}

// setInitial: MotionPlanner::MotionPlanner() (motion_planner.cpp:130-177); the velocity command is zero at construction
WBC_DEV void planner_init(const PlannerParams& P,PlannerState& s){
    for(int i=0;i<3;i++){ s.cmd[i]=0.0; s.vrot[i]=0.0; }
    s.yaw=0.0; s.step_time=0.0; s.cycle_time=0.0; s.step_phase=0; s.cycle_counter=0; s.mode=PL_IDLE; s.pad_=0;
    s.pi_body[0]=0.0; s.pi_body[1]=0.0; s.pi_body[2]=P.body_height;
    for(int i=0;i<3;i++) s.pf_body[i]=padd(s.pi_body[i],pmul(P.step_length,0.0));
    const double lhx=psub(s.pi_body[0],P.x_offset), lhy=padd(s.pi_body[1],P.y_offset);
    const double dx=pmul(2.0,P.x_offset), dy=-pmul(2.0,P.y_offset);
    // LH, LF = LH + (2x,0,0), RF = LH + (2x,-2y,0), RH = LH + (0,-2y,0)
    const double ox[4]={0.0,dx,dx,0.0}, oy[4]={0.0,0.0,dy,dy};
    for(int l=0;l<4;l++){ s.pi_foot[l][0]=(l==0)?lhx:padd(lhx,ox[l]); s.pi_foot[l][1]=(l==0)?lhy:padd(lhy,oy[l]); s.pi_foot[l][2]=(l==0)?0.0:padd(0.0,0.0);
                          for(int i=0;i<3;i++) s.pf_foot[l][i]=s.pi_foot[l][i]; }
    for(int i=0;i<6;i++){ s.pose[i]=0.0; s.vel[i]=0.0; s.acc[i]=0.0; }
    s.pose[2]=P.body_height;
    for(int i=0;i<12;i++){ s.swp[i]=0.0; s.swv[i]=0.0; s.swa[i]=0.0; }
    for(int l=0;l<4;l++){ s.contacts[l]=1; s.pad2_[l]=0; }
}

// one sample of the running step phase (motion_planner.cpp:229-358); k = step_phase
WBC_DEV void planner_sample(const PlannerParams& P,PlannerState& s){
    const int leg_of_phase[4]={0,3,1,2};                       // phases run LH, RH, LF, RF (:181); message slots LH 0, LF 3, RF 6, RH 9
    const int L=leg_of_phase[s.step_phase];
    const double t=s.step_time, sf=quintic(P.poly_foot,t), sd=quintic_d1(P.poly_foot,t), sdd=quintic_d2(P.poly_foot,t);
    const double oms=psub(1.0,sf), h=P.height_control_point;
    // Bernstein weights in the reference's evaluation order (:9-12, :22-25, :33-35)
    const double b0=pmul(pmul(oms,oms),oms), b1=pmul(pmul(pmul(3.0,oms),oms),sf), b2=pmul(pmul(pmul(3.0,oms),sf),sf), b3=pmul(pmul(sf,sf),sf);
    const double d0=pmul(oms,oms), d1=pmul(pmul(2.0,oms),sf), d2=pmul(sf,sf);
    for(int i=0;i<3;i++){
        const double v=(i==2)?h:0.0;
        const double p0=s.pi_foot[L][i], p1=padd(p0,v), p3=s.pf_foot[L][i], p2=padd(p3,v);
        double p=pmul(b0,p0); p=padd(p,pmul(b1,p1)); p=padd(p,pmul(b2,p2)); p=padd(p,pmul(b3,p3));
        const double f1=pmul(3.0,padd(padd(pmul(d0,psub(p1,p0)),pmul(d1,psub(p2,p1))),pmul(d2,psub(p3,p2))));
        const double f2=pmul(6.0,padd(pmul(oms,padd(psub(p2,pmul(2.0,p1)),p0)),pmul(sf,padd(psub(p3,pmul(2.0,p2)),p1))));
        s.swp[3*L+i]=p; s.swv[3*L+i]=pmul(f1,sd); s.swa[3*L+i]=padd(pmul(pmul(f2,sd),sd),pmul(f1,sdd));
    }
    for(int l=0;l<4;l++) s.contacts[l]=(l==L)?0:1;
    const double* pb=(s.cycle_counter==0)?P.poly_body_start:P.poly_body_continue;
    const double tc=s.cycle_time, sb=quintic(pb,tc), sbd=quintic_d1(pb,tc), sbdd=quintic_d2(pb,tc);
    for(int i=0;i<3;i++){ const double dl=psub(s.pf_body[i],s.pi_body[i]); s.pose[i]=padd(s.pi_body[i],pmul(sb,dl)); s.vel[i]=pmul(dl,sbd); s.acc[i]=pmul(dl,sbdd); }
    s.pose[3]=0.0; s.pose[4]=0.0; s.pose[5]=s.yaw; s.vel[3]=0.0; s.vel[4]=0.0; s.vel[5]=s.cmd[2]; s.acc[3]=0.0; s.acc[4]=0.0; s.acc[5]=0.0;
}

// one planner period; returns true if a message was published in it
WBC_DEV bool planner_tick(const PlannerParams& P,PlannerState& s){
    if(s.mode==PL_CYCLE_END){                                   // motion_planner.cpp:357-368
        s.cycle_counter++; s.step_phase=0; s.cycle_time=0.0;
        for(int i=0;i<3;i++){ s.pi_body[i]=s.pf_body[i]; s.pf_body[i]=padd(s.pf_body[i],pmul(s.vrot[i],P.step_length)); }
        for(int l=0;l<4;l++) for(int i=0;i<3;i++) s.pi_foot[l][i]=s.pf_foot[l][i];
        s.mode=PL_IDLE; return false;
    }
    if(s.mode==PL_IDLE){
        if(s.cmd[0]==0.0&&s.cmd[1]==0.0&&s.cmd[2]==0.0){        // :370-373 stand still
            for(int l=0;l<4;l++) s.contacts[l]=1;
            return true;
        }
        // start of a cycle (:201-222): rotate the command by the current yaw, advance the four foot targets
        const double cy=cos(s.yaw), sy=sin(s.yaw), vx=s.cmd[0], vy=s.cmd[1];
        s.vrot[0]=padd(padd(pmul(cy,vx),pmul(-sy,vy)),pmul(0.0,0.0)); s.vrot[1]=padd(padd(pmul(sy,vx),pmul(cy,vy)),pmul(0.0,0.0));
        s.vrot[2]=padd(padd(pmul(0.0,vx),pmul(0.0,vy)),pmul(1.0,0.0));
        const double dyaw=pmul(s.cmd[2],P.cycle_duration), cd=cos(dyaw), sdl=sin(dyaw);
        for(int l=0;l<4;l++){
            const double ax=psub(s.pi_foot[l][0],s.pi_body[0]), ay=psub(s.pi_foot[l][1],s.pi_body[1]);
            const double rot[3]={ padd(padd(pmul(cd,ax),pmul(-sdl,ay)),pmul(0.0,0.0)), padd(padd(pmul(sdl,ax),pmul(cd,ay)),pmul(0.0,0.0)),
                                  padd(padd(pmul(0.0,ax),pmul(0.0,ay)),pmul(1.0,0.0)) };
            const double vec[3]={ax,ay,0.0};
            for(int i=0;i<3;i++) s.pf_foot[l][i]=padd(s.pf_foot[l][i],padd(pmul(s.vrot[i],P.step_length),psub(rot[i],vec[i])));
        }
        s.mode=PL_IN_CYCLE;
    }
    // one pass of the inner loop (:224-356)
    if(s.step_time<P.step_duration){
        planner_sample(P,s);
        s.yaw=padd(s.yaw,pmul(s.cmd[2],P.dt)); s.step_time=padd(s.step_time,P.dt); s.cycle_time=padd(s.cycle_time,P.dt);
        return true;
    }
    s.step_phase++; s.step_time=0.0;
    if(s.step_phase==4) s.mode=PL_CYCLE_END;
    return false;
}

}  // namespace wbc
