#define DLBC_NC 3
#define DLBC_EXACT 0
#include "step_cN.inl"
