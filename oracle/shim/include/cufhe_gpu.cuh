// Header-swap used ONLY to build the reference's own test programs against this engine
// (oracle/Makefile target `reftests`): the reference tests say #include <include/cufhe_gpu.cuh>;
// with -Ioracle/shim ahead of -I/root/reference that name resolves here.
#pragma once
#include "../../../include/cufhe_b200.hpp"
