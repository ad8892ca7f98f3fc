#define DLBC_NC 1
#include "step_cN.inl"
