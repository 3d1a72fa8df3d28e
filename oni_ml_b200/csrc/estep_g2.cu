// E-step instantiations for G = 2 lanes per document (see estep_inst.inc).
#define ONI_G 2
#include "estep_inst.inc"
