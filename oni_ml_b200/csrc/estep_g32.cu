// E-step instantiations for G = 32 lanes per document (see estep_inst.inc).
#define ONI_G 32
#include "estep_inst.inc"
