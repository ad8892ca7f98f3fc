#define DLBC_NC 4
#include "step_cN.inl"
