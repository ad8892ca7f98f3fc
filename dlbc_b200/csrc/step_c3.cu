#define DLBC_NC 3
#include "step_cN.inl"
