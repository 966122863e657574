// See ref_order_shim.h. TEST INFRASTRUCTURE.
#include "objects/assembler.hpp"

bool std::less<Assembler*>::operator()(Assembler* a, Assembler* b) const { return a->id < b->id; }
