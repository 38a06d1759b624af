// Hex helpers the node uses around the PoW calls (reference src/crypto/ethash/helpers.hpp:12-54):
// to_hex(hash), to_hash256(hex), hash256 equality, and the shared epoch-0 context of the unit tests.
#pragma once
#include "ethash_compat.hpp"

#include <string>

template <typename Hash>
inline std::string to_hex(const Hash& h)
{
    static const char digits[] = "0123456789abcdef";
    std::string s(sizeof(h.bytes) * 2, '0');
    for (size_t i = 0; i < sizeof(h.bytes); ++i) {
        s[2 * i] = digits[h.bytes[i] >> 4];
        s[2 * i + 1] = digits[h.bytes[i] & 15];
    }
    return s;
}

inline ethash::hash256 to_hash256(const std::string& hex)
{
    const auto nibble = [](char c) -> int { return c <= '9' ? c - '0' : c - 'a' + 10; };
    ethash::hash256 h = {};
    for (size_t i = 0; i + 1 < hex.size() && i / 2 < sizeof(h.bytes); i += 2)
        h.bytes[i / 2] = static_cast<uint8_t>((nibble(hex[i]) << 4) | nibble(hex[i + 1]));
    return h;
}

inline bool operator==(const ethash::hash256& a, const ethash::hash256& b) noexcept { return std::memcmp(a.bytes, b.bytes, sizeof a) == 0; }
inline bool operator!=(const ethash::hash256& a, const ethash::hash256& b) noexcept { return !(a == b); }

inline const ethash::epoch_context& get_ethash_epoch_context_0() noexcept
{
    static ethash::epoch_context_ptr context = ethash::create_epoch_context(0);
    return *context;
}
