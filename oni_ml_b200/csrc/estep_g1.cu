// E-step instantiations for G = 1 lanes per document (see estep_inst.inc).
#define ONI_G 1
#include "estep_inst.inc"
