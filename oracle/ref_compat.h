// TEST INFRASTRUCTURE (oracle): force-included (-include) when compiling the UNMODIFIED reference
// sources from /root/reference into oracle/_ref/.  The reference does not compile on Linux/glibc as
// shipped: multi_special_primes.cpp:25 and :273 pass `uint64_t*` (unsigned long*) where
// util/uintarith.h:942-944 and :51 expect `unsigned long long*`.  Instead of patching
// the reference, this header adds two overloads that forward to the reference's own functions
// through a reinterpret_cast -- same object representation, no semantic change.
#pragma once
#include "seal/util/uintarith.h"
#include <cstddef>
#include <cstdint>

namespace seal
{
    namespace util
    {
        template <std::size_t Count>
        inline void multiply_accumulate_uint64(
            const std::uint64_t *operand1, const std::uint64_t *operand2, std::uint64_t *accumulator)
        {
            multiply_accumulate_uint64<Count>(
                operand1, operand2, reinterpret_cast<unsigned long long *>(accumulator));
        }

        inline unsigned char add_uint128(
            const std::uint64_t *operand1, const unsigned long long *operand2, std::uint64_t *result)
        {
            return add_uint128(operand1, operand2, reinterpret_cast<unsigned long long *>(result));
        }
    } // namespace util
} // namespace seal
