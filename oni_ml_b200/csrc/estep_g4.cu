// E-step instantiations for G = 4 lanes per document (see estep_inst.inc).
#define ONI_G 4
#include "estep_inst.inc"
