#define DLBC_NC 1
#define DLBC_EXACT 1
#include "step_cN.inl"
