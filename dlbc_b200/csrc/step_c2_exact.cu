#define DLBC_NC 2
#define DLBC_EXACT 1
#include "step_cN.inl"
