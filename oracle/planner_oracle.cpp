// ORACLE -- TEST INFRASTRUCTURE ONLY. See planner_oracle.hpp. Restates src/motion_planner.cpp of the reference.
#include "planner_oracle.hpp"
#include <cmath>

namespace oracle {
namespace {

struct P3 { double x=0,y=0,z=0; };
inline P3 operator+(const P3&a,const P3&b){ return {a.x+b.x,a.y+b.y,a.z+b.z}; }
inline P3 operator-(const P3&a,const P3&b){ return {a.x-b.x,a.y-b.y,a.z-b.z}; }
inline P3 operator*(double s,const P3&a){ return {s*a.x,s*a.y,s*a.z}; }
inline P3 operator*(const P3&a,double s){ return {a.x*s,a.y*s,a.z*s}; }
inline bool isZero(const P3&a){ return a.x==0.0&&a.y==0.0&&a.z==0.0; }
// 3x3 (row-major) times vector, coefficient-wise like a fixed-size Eigen product
inline P3 mul(const double* R,const P3&v){ return { R[0]*v.x+R[1]*v.y+R[2]*v.z, R[3]*v.x+R[4]*v.y+R[5]*v.z, R[6]*v.x+R[7]*v.y+R[8]*v.z }; }

// motion_planner.cpp:5-15
P3 bezier_4_points(double s,const P3&p0,const P3&p1,const P3&p2,const P3&p3){
    P3 p;
    p = (1-s)*(1-s)*(1-s) * p0;
    p = p + 3 * (1-s)*(1-s) * s * p1;
    p = p + 3 * (1-s) * s*s * p2;
    p = p + s*s*s * p3;
    return p;
}
// :18-26
P3 bezier_4_points_first_derivative(double s,const P3&p0,const P3&p1,const P3&p2,const P3&p3){
    double one_minus_s = 1.0 - s;
    P3 term1 = one_minus_s * one_minus_s * (p1 - p0);
    P3 term2 = 2.0 * one_minus_s * s * (p2 - p1);
    P3 term3 = s * s * (p3 - p2);
    return 3.0 * (term1 + term2 + term3);
}
// :29-36
P3 bezier_4_points_second_derivative(double s,const P3&p0,const P3&p1,const P3&p2,const P3&p3){
    double one_minus_s = 1.0 - s;
    P3 term1 = one_minus_s * (p2 - 2.0 * p1 + p0);
    P3 term2 = s * (p3 - 2.0 * p2 + p1);
    return 6.0 * (term1 + term2);
}
// :39-52: control points at a fixed height above the end points
P3 bezier(double s,const P3&pi,const P3&pf,double h){ P3 v{0.0,0.0,h}; return bezier_4_points(s,pi,pi+v,pf+v,pf); }
P3 bezier_first_derivative(double s,const P3&pi,const P3&pf,double h){ P3 v{0.0,0.0,h}; return bezier_4_points_first_derivative(s,pi,pi+v,pf+v,pf); }
P3 bezier_second_derivative(double s,const P3&pi,const P3&pf,double h){ P3 v{0.0,0.0,h}; return bezier_4_points_second_derivative(s,pi,pi+v,pf+v,pf); }
// :55-57
P3 segment(double s,const P3&pi,const P3&pf){ return pi + s*(pf-pi); }

// :59-97
struct QuinticPolynomial {
    double a0,a1,a2,a3,a4,a5;
    double evaluate(double t) const { return a0 + a1 * t + a2 * t * t + a3 * t * t * t + a4 * t * t * t * t + a5 * t * t * t * t * t; }
    double evaluateFirstDerivative(double t) const { return a1 + 2.0 * a2 * t + 3.0 * a3 * t * t + 4.0 * a4 * t * t * t + 5.0 * a5 * t * t * t * t; }
    double evaluateSecondDerivative(double t) const { return 2.0 * a2 + 6.0 * a3 * t + 12.0 * a4 * t * t + 20.0 * a5 * t * t * t; }
};
QuinticPolynomial generateQuinticPolynomial(double total_period,double vi=0.0,double vf=0.0){
    QuinticPolynomial poly; double T=total_period, T2=T*T, T3=T2*T, T4=T3*T, T5=T4*T;
    poly.a0 = 0.0; poly.a1 = vi; poly.a2 = 0.0;
    poly.a3 = (10.0 - 4.0 * vf * T - 6.0 * vi * T) / T3;
    poly.a4 = (-15.0 + 7.0 * vf * T + 8.0 * vi * T) / T4;
    poly.a5 = (6.0 - 3.0 * vf * T - 3.0 * vi * T) / T5;
    return poly;
}

}  // namespace

std::vector<PlannerRecord> planner_run(const PlannerParams& params,const std::vector<std::array<double,3>>& cmd){
    std::vector<PlannerRecord> out; const size_t K=cmd.size();
    if(K==0) return out;
    const double cycle_duration = 4 * params.step_duration;                         // :118
    // ---- constructor (:130-177); the velocity command is still zero here ----
    P3 velocity_command_; double yaw_rate_command_=0.0;
    ReferenceMsg ref_msg_{};
    ref_msg_.desiredComPose[2]=params.body_height;
    for(int i=0;i<4;i++) ref_msg_.footContacts[i]=1;
    double yaw_=0.0;
    P3 pi_body_{0.0,0.0,params.body_height};
    P3 pf_body_ = pi_body_ + params.step_length * velocity_command_;
    P3 RH_direction{0.0,-2*params.y_offset,0.0}, LF_direction{2*params.x_offset,0.0,0.0}, RF_direction{2*params.x_offset,-2*params.y_offset,0.0};
    P3 pi_foot_LH_{pi_body_.x-params.x_offset, pi_body_.y+params.y_offset, 0.0};
    P3 pi_foot_RH_ = pi_foot_LH_ + RH_direction, pi_foot_LF_ = pi_foot_LH_ + LF_direction, pi_foot_RF_ = pi_foot_LH_ + RF_direction;
    P3 pf_foot_LH_=pi_foot_LH_, pf_foot_RH_=pi_foot_RH_, pf_foot_LF_=pi_foot_LF_, pf_foot_RF_=pi_foot_RF_;

    // ---- pacing: one sleep() = end of one planner period ----
    size_t period=0; bool published=false;
    auto deliver=[&](){ velocity_command_.x=cmd[period][0]; velocity_command_.y=cmd[period][1]; yaw_rate_command_=cmd[period][2]; };   // input_callback :122-127
    auto publish=[&](){ published=true; };
    auto sleep=[&]()->bool{ out.push_back(PlannerRecord{published,ref_msg_}); published=false; period++; if(period>=K) return false; deliver(); return true; };
    deliver();

    // ---- plannerLoop (:180-376) ----
    int step_phase = 0, cycle_counter = 0; double current_step_time = 0.0, current_cycle_time = 0.0;
    QuinticPolynomial poly_foot = generateQuinticPolynomial(params.step_duration, 0.0, 0.0);
    QuinticPolynomial poly_body_start = generateQuinticPolynomial(cycle_duration, 0.0, params.body_final_velocity);
    QuinticPolynomial poly_body_continue = generateQuinticPolynomial(cycle_duration, params.body_final_velocity, params.body_final_velocity);
    while(true){
        if(!isZero(velocity_command_) || yaw_rate_command_!=0){
            double rot_matrix[9]={cos(yaw_),-sin(yaw_),0, sin(yaw_),cos(yaw_),0, 0,0,1};
            P3 velocity_command_rotated = mul(rot_matrix,velocity_command_);
            double delta_yaw = yaw_rate_command_ * cycle_duration;
            double rot_matrix_delta[9]={cos(delta_yaw),-sin(delta_yaw),0, sin(delta_yaw),cos(delta_yaw),0, 0,0,1};
            P3 LH_vec{pi_foot_LH_.x-pi_body_.x, pi_foot_LH_.y-pi_body_.y, 0.0}, RH_vec{pi_foot_RH_.x-pi_body_.x, pi_foot_RH_.y-pi_body_.y, 0.0};
            P3 LF_vec{pi_foot_LF_.x-pi_body_.x, pi_foot_LF_.y-pi_body_.y, 0.0}, RF_vec{pi_foot_RF_.x-pi_body_.x, pi_foot_RF_.y-pi_body_.y, 0.0};
            P3 direction_new_LH = mul(rot_matrix_delta,LH_vec) - LH_vec, direction_new_RH = mul(rot_matrix_delta,RH_vec) - RH_vec;
            P3 direction_new_LF = mul(rot_matrix_delta,LF_vec) - LF_vec, direction_new_RF = mul(rot_matrix_delta,RF_vec) - RF_vec;
            pf_foot_LH_ = pf_foot_LH_ + (velocity_command_rotated * params.step_length + direction_new_LH);
            pf_foot_RH_ = pf_foot_RH_ + (velocity_command_rotated * params.step_length + direction_new_RH);
            pf_foot_LF_ = pf_foot_LF_ + (velocity_command_rotated * params.step_length + direction_new_LF);
            pf_foot_RF_ = pf_foot_RF_ + (velocity_command_rotated * params.step_length + direction_new_RF);
            while(step_phase < 4){
                if(current_step_time < params.step_duration){
                    double s_foot = poly_foot.evaluate(current_step_time), s_foot_dot = poly_foot.evaluateFirstDerivative(current_step_time),
                           s_foot_ddot = poly_foot.evaluateSecondDerivative(current_step_time);
                    const P3* pi[4]={&pi_foot_LH_,&pi_foot_RH_,&pi_foot_LF_,&pi_foot_RF_}; const P3* pf[4]={&pf_foot_LH_,&pf_foot_RH_,&pf_foot_LF_,&pf_foot_RF_};
                    static const int slot[4]={0,9,3,6};                                        // message offsets of LH, RH, LF, RF (:253-310)
                    static const unsigned char contacts[4][4]={{0,1,1,1},{1,1,1,0},{1,0,1,1},{1,1,0,1}};
                    if(step_phase>=0&&step_phase<4){ const int k=step_phase; const double h=params.height_control_point;
                        P3 point = bezier(s_foot,*pi[k],*pf[k],h);
                        P3 point_dot = bezier_first_derivative(s_foot,*pi[k],*pf[k],h) * s_foot_dot;
                        P3 point_ddot = bezier_second_derivative(s_foot,*pi[k],*pf[k],h) * s_foot_dot * s_foot_dot
                                      + bezier_first_derivative(s_foot,*pi[k],*pf[k],h) * s_foot_ddot;
                        double* P=ref_msg_.desiredSwingLegsPosition+slot[k]; double* V=ref_msg_.desiredSwingLegsVelocity+slot[k]; double* A=ref_msg_.desiredSwingLegsAcceleration+slot[k];
                        P[0]=point.x; P[1]=point.y; P[2]=point.z; V[0]=point_dot.x; V[1]=point_dot.y; V[2]=point_dot.z; A[0]=point_ddot.x; A[1]=point_ddot.y; A[2]=point_ddot.z;
                        for(int i=0;i<4;i++) ref_msg_.footContacts[i]=contacts[k][i];
                    }
                    double s_body,s_body_dot,s_body_ddot;
                    if(cycle_counter == 0){ s_body=poly_body_start.evaluate(current_cycle_time); s_body_dot=poly_body_start.evaluateFirstDerivative(current_cycle_time);
                                            s_body_ddot=poly_body_start.evaluateSecondDerivative(current_cycle_time); }
                    else { s_body=poly_body_continue.evaluate(current_cycle_time); s_body_dot=poly_body_continue.evaluateFirstDerivative(current_cycle_time);
                           s_body_ddot=poly_body_continue.evaluateSecondDerivative(current_cycle_time); }
                    P3 point_segment_body = segment(s_body,pi_body_,pf_body_);
                    P3 point_segment_body_dot = (pf_body_ - pi_body_) * s_body_dot, point_segment_body_ddot = (pf_body_ - pi_body_) * s_body_ddot;
                    double* D=ref_msg_.desiredComPose; D[0]=point_segment_body.x; D[1]=point_segment_body.y; D[2]=point_segment_body.z; D[3]=0.0; D[4]=0.0; D[5]=yaw_;
                    double* W=ref_msg_.desiredComVelocity; W[0]=point_segment_body_dot.x; W[1]=point_segment_body_dot.y; W[2]=point_segment_body_dot.z; W[3]=0.0; W[4]=0.0; W[5]=yaw_rate_command_;
                    double* C=ref_msg_.desiredComAcceleration; C[0]=point_segment_body_ddot.x; C[1]=point_segment_body_ddot.y; C[2]=point_segment_body_ddot.z; C[3]=0.0; C[4]=0.0; C[5]=0.0;
                    publish();
                    yaw_ += yaw_rate_command_ * params.dt;
                    current_step_time += params.dt;
                    current_cycle_time += params.dt;
                } else {
                    step_phase = step_phase + 1;
                    current_step_time = 0.0;
                }
                if(!sleep()) return out;
            }
            if(step_phase == 4){
                cycle_counter++; step_phase = 0; current_cycle_time = 0.0;
                pi_body_ = pf_body_;
                pf_body_ = pf_body_ + velocity_command_rotated*params.step_length;
                pi_foot_LH_=pf_foot_LH_; pi_foot_RH_=pf_foot_RH_; pi_foot_LF_=pf_foot_LF_; pi_foot_RF_=pf_foot_RF_;
            }
        } else {
            for(int i=0;i<4;i++) ref_msg_.footContacts[i]=1;
            publish();
        }
        if(!sleep()) return out;
    }
}

}  // namespace oracle
