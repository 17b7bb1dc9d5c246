/* Using libddpb200.so from plain C: the host-side planner needs no GPU, so this also runs on a build box.
 *
 *   gcc -std=c99 -Iinclude examples/c_api_plan.c -o /tmp/c_api_plan ddpb200/_lib/libddpb200.so -Wl,-rpath,$PWD/ddpb200/_lib
 *   /tmp/c_api_plan [batch]
 *
 * Prints, for every VGG-11 layer, the (tile width, split-K) plan of the fused conv + BatchNorm forward and of dgrad,
 * and the scratch / sync-word sizes a caller has to provide ("From C / C++").
 */
#include <stdio.h>
#include <stdlib.h>

#include "ddpb200.h"

int main(int argc, char** argv) {
  const int batch = argc > 1 ? atoi(argv[1]) : 256;
  static const int layers[8][4] = {/* H = W, C_in of the GEMM, C_out, taps */
                                   {32, 64, 64, 1},  {16, 64, 128, 9}, {8, 128, 256, 9}, {8, 256, 256, 9},
                                   {4, 256, 512, 9}, {4, 512, 512, 9}, {2, 512, 512, 9}, {2, 512, 512, 9}};
  printf("libddpb200 ABI %d, planning for %d SMs, batch %d, %d sync words per stream\n", ddpb_abi_version(),
         ddpb_sm_count(), batch, ddpb_conv_split_sync_words());
  for (int l = 0; l < 8; ++l) {
    const int s = layers[l][0], ci = layers[l][1], co = layers[l][2], taps = layers[l][3];
    int fwd[4], dg[4] = {0, 0, 0, 0};
    ddpb_conv_plan_ex(batch, s, s, ci, co, taps, 0, /*with_tail=*/1, /*with_sync=*/1, fwd);
    if (l > 0) ddpb_conv_plan_ex(batch, s, s, co, ci, 9, 0, 0, 0, dg);
    printf("L%d %2dx%-2d %3d->%3d | fwd: %3d-wide tile x%d%s, scratch %lld floats", l, s, s, ci, co, fwd[1], fwd[2],
           fwd[3] ? " (summed in the kernel)" : "", ddpb_conv_fprop_scratch_elems(batch, s, s, ci, co, taps, 0));
    if (l > 0) printf(" | dgrad: %3d-wide x%d", dg[1], dg[2]);
    printf(" | wgrad: %d pixel splits, scratch %lld floats\n", ddpb_conv_wgrad_splits(batch, s, s, ci, co, taps),
           ddpb_conv_wgrad_scratch_elems(batch, s, s, ci, co, taps));
  }
  return 0;
}
