// k_pcg_g16.cu — the second build of the persistent PCG kernel (k_pcg.cu): ring geometry 16 warps x 2 stages x 16 slots
// instead of 20 x 2 x 12. A slice of the 2-D PPE has ~28 slots: with 16-slot stages both of its stages are in flight
// at once, so a warp finishes a slice in one TMA round trip less. That shortens the tail of sweep A and wins where the
// synchronisations weigh most — row blocks of up to ~1.3 M rows (a rank's block on 4 or 8 GPUs, mid-size single-GPU
// cases); at 4 M rows on one GPU the 20 x 2 x 12 ring streams 3.5 % faster and stays. Same code, same results.
#undef PCG_NWARPS
#undef PCG_NCH
#undef PCG_NSTAGES
#define PCG_NWARPS 16
#define PCG_NCH 16
#define PCG_NSTAGES 2
#define PCG_VARIANT 1
#include "k_pcg.cu"
