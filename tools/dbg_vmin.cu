// Reproducer for an nvcc 12.9 / sm_100a miscompile: max(max(a, -b), ...) fused into VIMNMX3.RELU
// reads the un-negated register.  Expected amax 32555; the max(a,-b) form printed 5 on a B200.
#include <cstdio>
#include <cstdint>
__global__ void k(const uint32_t* in, int n, uint32_t* out){
  uint32_t mx=0x80008000u, mn=0x7FFF7FFFu;
  for(int i=0;i<n;i++){ uint32_t w=in[i]; mx=__vmaxs2(mx,w); mn=__vmins2(mn,w);}  
  out[0]=mx; out[1]=mn;
  int hi = max((int)mx >> 16, -((int)mn >> 16));
  int lo = max((int)(int16_t)(mx & 0xFFFFu), -(int)(int16_t)(mn & 0xFFFFu));
  out[2]=max(max(hi,lo),0);
  { int a=abs((int)mx>>16), b=abs((int)mn>>16), c=abs((int)(int16_t)(mx&0xFFFFu)), d=abs((int)(int16_t)(mn&0xFFFFu)); out[5]=max(max(a,b),max(c,d)); }
  int mnh=(int)(int16_t)(mn>>16); out[3]=(uint32_t)(-mnh);
  out[4]=(uint32_t)(-((int)mn >> 16));
}
int main(){
  uint32_t h[16]; for(int i=0;i<16;i++) h[i]=0x00050005u;
  int16_t v=-32555; h[3]=(uint32_t)(uint16_t)5 | ((uint32_t)(uint16_t)v<<16);
  uint32_t *d,*o; cudaMalloc(&d,64); cudaMalloc(&o,64); cudaMemcpy(d,h,64,cudaMemcpyHostToDevice);
  k<<<1,1>>>(d,16,o); uint32_t r[6]; cudaMemcpy(r,o,24,cudaMemcpyDeviceToHost);
  printf("mx=%08x mn=%08x amax_buggy=%d negmnh=%d negmnh2=%d amax_abs=%d err=%s\n",r[0],r[1],(int)r[2],(int)r[3],(int)r[4],(int)r[5],cudaGetErrorString(cudaGetLastError()));
}
