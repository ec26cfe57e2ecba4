// disk_cache.hpp -- the one on-disk cache of the library: verified blobs under TFK_CACHE_DIR (default
// ~/.cache/tinyfk_b200).  Two users: the plan-specialised cubins (jit.cpp) and the flat link / joint tables of a parsed
// URDF (model.cpp; SURVEY.md section 8(f) rank 4, "URDF -> flat-table cache on disk").
//
// A cache file = header + payload.  The header carries a magic, two independent 64-bit hashes of everything that
// determines the payload (the caller's key), the payload length and a hash of the payload: a truncated, foreign,
// colliding or older-layout file is detected, deleted and rebuilt instead of being trusted.  Only a directory that
// belongs to this user and that nobody else can write to is used; files are written through a temporary and renamed,
// so concurrent processes never see a partial file.  Everything here is best effort: any failure means "no cache".
#ifndef TFK_DISK_CACHE_HPP
#define TFK_DISK_CACHE_HPP

#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>

namespace tfk {
namespace diskcache {

uint64_t fnv1a_bytes(const void *data, size_t n, uint64_t h = 1469598103934665603ull);
inline uint64_t fnv1a(const std::string &s, uint64_t h = 1469598103934665603ull) { return fnv1a_bytes(s.data(), s.size(), h); }
const uint64_t k_second_offset = 0x9e3779b97f4a7c15ull; // start value of the second hash stream

// the trusted cache directory (created when missing), or "" when there is none
std::string dir();

// reads <dir>/<file>; true and the payload when the header matches (magic, both keys, length, payload hash); a file
// that does not match is removed
bool read(const std::string &file, const char magic[8], uint64_t key_a, uint64_t key_b, std::vector<char> &payload,
          size_t min_payload = 1);
void write(const std::string &file, const char magic[8], uint64_t key_a, uint64_t key_b, const char *data, size_t n);

} // namespace diskcache
} // namespace tfk
#endif
