/* CPU check of the divide-free cell index of csrc/bvx_kernels.cuh (cell_of): for every axis value whose fast
 * path is taken (t = (v - lo) * RN(1/size) further than the margin 2^-21 (grid + 1) from both neighbouring
 * integers), floor(t) must equal floor(RN((v - lo) / size)), or both must lie outside the grid.  Plain IEEE fp32,
 * compile with -ffp-contract=off.  usage: cell_axis_check [iterations per (size, origin) pair] */
#include <stdio.h>
#include <stdlib.h>
#include <math.h>
#include <stdint.h>
#include <string.h>
static uint64_t s = 88172645463325252ull;
static uint64_t rnd(){ s ^= s << 13; s ^= s >> 7; s ^= s << 17; return s; }
int main(int argc, char** argv){
  long iters = argc > 1 ? atol(argv[1]) : 3000000;
  float sizes[] = {0.05f,0.1f,0.16f,4.0f,0.2f,0.075f,0.32f,8.0f,0.25f,1.0f/3.0f,0.0123f,1e-3f,7.0f};
  float los[] = {0.f,-40.f,-3.f,-39.68f,-51.2f,-5.f,-75.2f,-2.f, 0.1f};
  float exts[] = {70.4f, 80.f, 4.f, 79.36f, 102.4f, 8.f, 150.4f, 6.f, 33.3f};
  long bad=0, total=0, slow=0;
  for (int si=0; si<13; ++si) for (int li=0; li<9; ++li){
    float vs=sizes[si], lo=los[li]; volatile float r = 1.0f/vs; float rvs=r;
    volatile float q0 = exts[li]/vs; long gs = (long)rintf(q0); if (gs < 1) gs = 1; if (gs > (1<<24)) continue;
    float gsf = (float)gs; float mg = nextafterf((float)(((double)gs+1.0)/2097152.0), INFINITY);
    for (long it=0; it<iters; ++it){
      float v; uint64_t x = rnd(); int mode = x & 7;
      if (mode < 4) { v = lo + (float)((x>>8)%400000)/1000.0f*exts[li]/200.0f - exts[li]*0.5f; }
      else if (mode < 6) { long k = (long)((x>>8)%(gs+200)) - 100; float b = lo + k*vs; uint32_t u; memcpy(&u,&b,4); u += (int)((x>>40)%9) - 4; memcpy(&v,&u,4); }
      else { uint32_t u = (uint32_t)(x>>16); memcpy(&v,&u,4); }
      volatile float a = v - lo; volatile float t = a * rvs; float k = floorf(t); volatile float fr = t - k; volatile float omf = 1.0f - fr;
      int safe = (fr > mg) & (omf > mg);
      volatile float qq = a / vs; float kr = floorf(qq);
      total++;
      if (!safe) { slow++; continue; }
      int in_f = (k >= 0.0f) && (k < gsf), in_r = (kr >= 0.0f) && (kr < gsf);
      if (in_f != in_r || (in_f && k != kr)) { if (bad<10) printf("BAD v=%a lo=%a vs=%a gs=%ld k=%a kr=%a\n",v,lo,vs,gs,k,kr); bad++; }
    }
  }
  printf("total %ld bad %ld slow %ld\n", total, bad, slow);
  return bad != 0;
}
