/* The rope models built on mb_chain.h: name, links, flexible?  One list for the CUDA kernel set
 * (csrc/mb_model_chain.cu), the oracle (oracle/mbo_model_chain.c) and, by name, the Python host (mechanics/chain.py).
 * Geometry follows the generated models of mechanics/zoo.py: stiff `Rope(l=1, m=6, dir=[0.05,1,0])`
 * (docs/src/examples/ropes_and_cables.md:24-36), flexible `Rope(l=1, m=5, c, d, dir=[0.2,1,0])` (:48-60). */
#ifndef MB_CHAIN_MODELS_H
#define MB_CHAIN_MODELS_H
#define MBC_MAX_LINKS 50
#define MBC_CHAIN_MODELS(X) \
  X(rope6c, 6, 0) X(rope3ec, 3, 1) X(rope20, 20, 0) X(rope50, 50, 0) X(rope20e, 20, 1) X(rope50e, 50, 1)
MBC_FN void mbc_rope_init(MbcRope* R, int N, int elastic) {
  const double dx = elastic ? 0.2 : 0.05, nn = sqrt(dx * dx + 1.0);
  R->nlinks = N; R->elastic = elastic;
  R->dir[0] = dx / nn; R->dir[1] = 1.0 / nn; R->dir[2] = 0.0;
  R->l = 1.0; R->m = elastic ? 5.0 : 6.0;
}
#endif
