#include <cuda_runtime.h>
__device__ __forceinline__ float rsq(float x){ float y; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
// variant A: pack across j (xi duplicated)
extern "C" __global__ void __launch_bounds__(256) kA(const float4* __restrict__ q, float4* out, int nq, float eps2)
{
    extern __shared__ float4 sm[];
    int t = blockIdx.x*blockDim.x + threadIdx.x;
    float4 X=q[t*4+0],Y=q[t*4+1],Z=q[t*4+2];
    float2 nx[4]={{-X.x,-X.x},{-X.y,-X.y},{-X.z,-X.z},{-X.w,-X.w}};
    float2 ny[4]={{-Y.x,-Y.x},{-Y.y,-Y.y},{-Y.z,-Y.z},{-Y.w,-Y.w}};
    float2 nz[4]={{-Z.x,-Z.x},{-Z.y,-Z.y},{-Z.z,-Z.z},{-Z.w,-Z.w}};
    float2 ax[4],ay[4],az[4];
    #pragma unroll
    for(int r=0;r<4;r++){ax[r]=make_float2(0,0);ay[r]=ax[r];az[r]=ax[r];}
    float2 e2=make_float2(eps2,eps2);
    for(int i=threadIdx.x;i<nq*4;i+=blockDim.x) sm[i]=q[i];
    __syncthreads();
    #pragma unroll 1
    for(int jq=0;jq<nq;jq++){
        float4 JX=sm[jq*4],JY=sm[jq*4+1],JZ=sm[jq*4+2],JM=sm[jq*4+3];
        #pragma unroll
        for(int p=0;p<2;p++){
            float2 xj=p?make_float2(JX.z,JX.w):make_float2(JX.x,JX.y);
            float2 yj=p?make_float2(JY.z,JY.w):make_float2(JY.x,JY.y);
            float2 zj=p?make_float2(JZ.z,JZ.w):make_float2(JZ.x,JZ.y);
            float2 mj=p?make_float2(JM.z,JM.w):make_float2(JM.x,JM.y);
            #pragma unroll
            for(int r=0;r<4;r++){
                float2 dx=__fadd2_rn(xj,nx[r]),dy=__fadd2_rn(yj,ny[r]),dz=__fadd2_rn(zj,nz[r]);
                float2 r2=__ffma2_rn(dx,dx,e2); r2=__ffma2_rn(dy,dy,r2); r2=__ffma2_rn(dz,dz,r2);
                float2 ri=make_float2(rsq(r2.x),rsq(r2.y));
                float2 ri2=__fmul2_rn(ri,ri); float2 mr=__fmul2_rn(mj,ri); float2 s=__fmul2_rn(ri2,mr);
                ax[r]=__ffma2_rn(dx,s,ax[r]); ay[r]=__ffma2_rn(dy,s,ay[r]); az[r]=__ffma2_rn(dz,s,az[r]);
            }
        }
    }
    out[t*3+0]=make_float4(ax[0].x+ax[0].y,ax[1].x+ax[1].y,ax[2].x+ax[2].y,ax[3].x+ax[3].y);
    out[t*3+1]=make_float4(ay[0].x+ay[0].y,ay[1].x+ay[1].y,ay[2].x+ay[2].y,ay[3].x+ay[3].y);
    out[t*3+2]=make_float4(az[0].x+az[0].y,az[1].x+az[1].y,az[2].x+az[2].y,az[3].x+az[3].y);
}
