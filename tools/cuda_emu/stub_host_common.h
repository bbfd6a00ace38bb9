// Emulation stand-in for csrc/host_common.h (no driver, no tensor maps).
#pragma once
#include <stdint.h>
#include "emu_cuda.h"
namespace omldm {
int set_error(const char* what, cudaError_t e);
int set_error_msg(const char* msg);
int check_launch(const char* kernel_name);
void count_launch(int n = 1);
}  // namespace omldm

// timeline stamps of tuning builds (csrc/host_common.h): nothing on the emulator
#define OMLDM_TL_DECLARE(name)
#define OMLDM_TL_BEGIN(name) do { } while (0)
#define OMLDM_TL_END(name) do { } while (0)
