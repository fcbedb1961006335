// dqmcb/detail.h -- glue between the C ABI's status codes and the reference's error behaviour.
#pragma once
#include <stdexcept>
#include <string>

#include "../dqmc_b200.h"

namespace dqmcb {

// C-ABI status -> the exception type the reference throws in the corresponding situation:
// std::invalid_argument / std::out_of_range for bad arguments, std::runtime_error for numerical
// failure (source/stablelinalg.cpp:43-45) and for CUDA failures.  There is no CPU fallback.
inline void check(int rc, const dqmcb_ctx* ctx, const char* what) {
    if (rc == DQMCB_OK) return;
    const char* msg = dqmcb_last_error(ctx);
    const std::string text = std::string(what) + ": " + (msg ? msg : "");
    switch (rc) {
        case DQMCB_EINVAL: throw std::invalid_argument(text);
        case DQMCB_ENUMERIC: throw std::runtime_error(text + " (factorisation failed)");
        case DQMCB_ESTATE: throw std::logic_error(text);
        default: throw std::runtime_error(text);
    }
}

}  // namespace dqmcb
