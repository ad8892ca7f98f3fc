#define DLBC_NC 3
#define DLBC_EXACT 1
#include "step_cN.inl"
