#include "lj_internal.h"
int lj_render_wavefront(lj_scene *s, const RenderLaunch &rl, double *dev_sum, lj_stats *st) {
    return lj_render_megakernel(s, rl, dev_sum, st);
}
