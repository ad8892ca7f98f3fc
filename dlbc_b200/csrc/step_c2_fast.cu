#define DLBC_NC 2
#define DLBC_EXACT 0
#include "step_cN.inl"
