#define DLBC_NC 4
#define DLBC_EXACT 0
#include "step_cN.inl"
