#include <cstdio>
#include <cstdlib>
#include <algorithm>
__device__ __host__ inline int ref_strength(const int* d){ int best=0; for(int k=0;k<16;k++){int mn=d[k],mx=d[k]; for(int j=1;j<9;j++){int v=d[(k+j)&15]; mn=mn<v?mn:v; mx=mx>v?mx:v;} best=best>mn?best:mn; best=best>-mx?best:-mx;} return best;}
__device__ int vA(const int* dd){ int d[16]; for(int k=0;k<16;k++) d[k]=dd[k];
    int mn3[16], mx3[16];
#pragma unroll
    for (int k = 0; k < 16; k++) { mn3[k] = __vimin3_s32(d[k], d[(k + 1) & 15], d[(k + 2) & 15]); mx3[k] = __vimax3_s32(d[k], d[(k + 1) & 15], d[(k + 2) & 15]); }
    int best = 0;
#pragma unroll
    for (int k = 0; k < 16; k++) { const int mn9 = __vimin3_s32(mn3[k], mn3[(k + 3) & 15], mn3[(k + 6) & 15]); const int mx9 = __vimax3_s32(mx3[k], mx3[(k + 3) & 15], mx3[(k + 6) & 15]); best = __vimax3_s32(best, mn9, -mx9); }
    return best; }
__device__ int vB(const int* dd){ int d[16]; for(int k=0;k<16;k++) d[k]=dd[k];
    int mn3[16], mx3[16];
#pragma unroll
    for (int k = 0; k < 16; k++) { mn3[k] = min(min(d[k], d[(k + 1) & 15]), d[(k + 2) & 15]); mx3[k] = max(max(d[k], d[(k + 1) & 15]), d[(k + 2) & 15]); }
    int best = 0;
#pragma unroll
    for (int k = 0; k < 16; k++) { const int mn9 = min(min(mn3[k], mn3[(k + 3) & 15]), mn3[(k + 6) & 15]); const int mx9 = max(max(mx3[k], mx3[(k + 3) & 15]), mx3[(k + 6) & 15]); best = max(max(best, mn9), -mx9); }
    return best; }
__device__ int vC(const int* dd){ int d[16]; for(int k=0;k<16;k++) d[k]=dd[k];
    int mn3[16], mx3[16];
#pragma unroll
    for (int k = 0; k < 16; k++) { mn3[k] = __vimin3_s32(d[k], d[(k + 1) & 15], d[(k + 2) & 15]); mx3[k] = __vimax3_s32(d[k], d[(k + 1) & 15], d[(k + 2) & 15]); }
    int bmn = 0, bmx = 0x7fffffff;
#pragma unroll
    for (int k = 0; k < 16; k++) { const int mn9 = __vimin3_s32(mn3[k], mn3[(k + 3) & 15], mn3[(k + 6) & 15]); const int mx9 = __vimax3_s32(mx3[k], mx3[(k + 3) & 15], mx3[(k + 6) & 15]); bmn = max(bmn, mn9); bmx = min(bmx, mx9); }
    return max(bmn, -bmx); }
__global__ void k(const int* d, int n, int* o){ int i=blockIdx.x*blockDim.x+threadIdx.x; if(i<n){ o[4*i]=ref_strength(d+16*i); o[4*i+1]=vA(d+16*i); o[4*i+2]=vB(d+16*i); o[4*i+3]=vC(d+16*i);} }
int main(){ int n=100000; int* h=(int*)malloc(n*16*4); srand(1); for(int i=0;i<n*16;i++) h[i]=(rand()%511)-255; 
 for(int i=0;i<n/2;i++){ int s=rand()%16, sg=(rand()&1)?1:-1; for(int j=0;j<10;j++) h[16*i+((s+j)&15)] = sg*(8+rand()%100);} 
 int *d,*o; cudaMalloc(&d,n*64); cudaMalloc(&o,n*16); cudaMemcpy(d,h,n*64,cudaMemcpyHostToDevice); k<<<(n+255)/256,256>>>(d,n,o); int* ho=(int*)malloc(n*16); cudaMemcpy(ho,o,n*16,cudaMemcpyDeviceToHost);
 int bad[4]={0,0,0,0}; for(int i=0;i<n;i++){ int r=ref_strength(h+16*i); for(int v=0;v<4;v++) if(ho[4*i+v]!=r) bad[v]++; } printf("mismatch devref=%d A=%d B=%d C=%d\n",bad[0],bad[1],bad[2],bad[3]); return 0; }
