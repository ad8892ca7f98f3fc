#define DLBC_NC 2
#include "step_cN.inl"
