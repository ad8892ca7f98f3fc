#define DLBC_NC 1
#define DLBC_EXACT 0
#include "step_cN.inl"
