// ref_order_shim.h -- TEST INFRASTRUCTURE, force-included (-include) into every translation unit of the
// reference when building oracle/_ref_ordered (oracle/build_ref_ordered.sh). Nothing else uses it.
//
// The reference's Clipper keeps std::set<Assembler*> (systems/clipper.hpp:24-26) and walks them in
// pick_assembler_to_clip (:207-229): the iteration order is the order of heap addresses, which differs from
// process to process and from env to env (measured), so *which* assembler a given random
// number selects cannot be reproduced. This shim replaces the comparator of exactly those sets,
// std::less<Assembler*>, by "lower object id first" (ids are given in creation order, core/grid.hpp:56-68).
// Every line of the reference's own clipper code still runs; only the unreproducible tie to malloc is gone.
// The simulator and the oracle use the same creation order, so this build pins their clipper against the
// reference's code (tests/golden/make_golden.py clipper -> tests/golden/clipper_spread.npz).
#ifndef ORACLE_REF_ORDER_SHIM_H_
#define ORACLE_REF_ORDER_SHIM_H_
#ifdef __cplusplus
#include <functional>

class Assembler;
namespace std {
template <>
struct less<Assembler*> {
  bool operator()(Assembler* a, Assembler* b) const;  // defined in ref_order_shim.cpp, where Assembler is complete
};
}  // namespace std
#endif
#endif  // ORACLE_REF_ORDER_SHIM_H_
