/*
 * k_chain.cu -- type dispatch for the fused typecast + arithmetic chain kernels.
 */
#include "k_chain.cuh"

namespace nnsb {

chain_fn chain_row_0 (int, bool);
chain_fn chain_row_1 (int, bool);
chain_fn chain_row_2 (int, bool);
chain_fn chain_row_3 (int, bool);
chain_fn chain_row_4 (int, bool);
chain_fn chain_row_5 (int, bool);
chain_fn chain_row_6 (int, bool);
chain_fn chain_row_7 (int, bool);
chain_fn chain_row_8 (int, bool);
chain_fn chain_row_9 (int, bool);
chain_fn chain_row_10 (int, bool);

int launch_chain (int in_type, int out_type, bool orc, const void *in, void *out, size_t n,
    const ChainArgs &args, const LaunchCfg &cfg, cudaStream_t stream)
{
  static chain_fn (*const rows[NNSB_TYPE_END]) (int, bool) = {
    chain_row_0, chain_row_1, chain_row_2, chain_row_3, chain_row_4, chain_row_5,
    chain_row_6, chain_row_7, chain_row_8, chain_row_9, chain_row_10
  };
  if (in_type < 0 || in_type >= NNSB_TYPE_END)
    return (int) cudaErrorInvalidValue;
  chain_fn fn = rows[in_type] (out_type, orc);
  if (!fn)
    return (int) cudaErrorInvalidValue;
  return fn (in, out, n, args, cfg, stream);
}

}  // namespace nnsb
