#define DLBC_NC 4
#define DLBC_EXACT 1
#include "step_cN.inl"
