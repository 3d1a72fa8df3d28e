// E-step instantiations for G = 16 lanes per document (see estep_inst.inc).
#define ONI_G 16
#include "estep_inst.inc"
