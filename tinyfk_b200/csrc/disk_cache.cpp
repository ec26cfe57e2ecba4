// disk_cache.cpp -- see disk_cache.hpp.
#include "disk_cache.hpp"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iterator>
#include <sys/stat.h>
#include <unistd.h>

namespace tfk {
namespace diskcache {
namespace {

struct Header {
  char magic[8];
  uint64_t key_a, key_b, payload_bytes, payload_hash;
};

std::string configured_dir() {
  if (const char *d = std::getenv("TFK_CACHE_DIR")) return *d ? std::string(d) : std::string();
  if (const char *h = std::getenv("HOME")) return std::string(h) + "/.cache/tinyfk_b200";
  return std::string();
}

void mkdirs(const std::string &p) {
  for (size_t i = 1; i <= p.size(); ++i)
    if (i == p.size() || p[i] == '/') mkdir(p.substr(0, i).c_str(), 0700);
}

// only a directory that belongs to this user and that nobody else can write to is trusted
bool trusted(const std::string &d) {
  struct stat st;
  if (stat(d.c_str(), &st) != 0) return false;
  return S_ISDIR(st.st_mode) && st.st_uid == geteuid() && (st.st_mode & (S_IWGRP | S_IWOTH)) == 0;
}

} // namespace

uint64_t fnv1a_bytes(const void *data, size_t n, uint64_t h) {
  const unsigned char *p = static_cast<const unsigned char *>(data);
  for (size_t i = 0; i < n; ++i) {
    h ^= p[i];
    h *= 1099511628211ull;
  }
  return h;
}

std::string dir() {
  std::string d = configured_dir();
  if (d.empty()) return d;
  mkdirs(d);
  if (!trusted(d)) d.clear(); // someone else's (or a writable-by-others) directory: no disk cache
  return d;
}

bool read(const std::string &file, const char magic[8], uint64_t key_a, uint64_t key_b, std::vector<char> &payload,
          size_t min_payload) {
  const std::string d = dir();
  if (d.empty()) return false;
  const std::string path = d + "/" + file;
  std::ifstream f(path, std::ios::binary);
  if (!f) return false;
  std::vector<char> all((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
  f.close();
  Header hd;
  bool ok = all.size() >= sizeof hd + min_payload;
  if (ok) {
    std::memcpy(&hd, all.data(), sizeof hd);
    ok = std::memcmp(hd.magic, magic, 8) == 0 && hd.key_a == key_a && hd.key_b == key_b &&
         hd.payload_bytes == all.size() - sizeof hd && hd.payload_hash == fnv1a_bytes(all.data() + sizeof hd, all.size() - sizeof hd);
  }
  if (!ok) {
    std::remove(path.c_str()); // corrupt, truncated, foreign or from an older layout: the caller rebuilds it
    return false;
  }
  payload.assign(all.begin() + sizeof hd, all.end());
  return true;
}

void write(const std::string &file, const char magic[8], uint64_t key_a, uint64_t key_b, const char *data, size_t n) {
  const std::string d = dir();
  if (d.empty()) return;
  const std::string path = d + "/" + file, tmp = path + "." + std::to_string((long)getpid()) + ".tmp";
  std::ofstream f(tmp, std::ios::binary);
  if (!f) return;
  Header hd;
  std::memcpy(hd.magic, magic, 8);
  hd.key_a = key_a;
  hd.key_b = key_b;
  hd.payload_bytes = n;
  hd.payload_hash = fnv1a_bytes(data, n);
  f.write(reinterpret_cast<const char *>(&hd), sizeof hd);
  f.write(data, (std::streamsize)n);
  f.close();
  if (!f || std::rename(tmp.c_str(), path.c_str()) != 0) std::remove(tmp.c_str());
}

} // namespace diskcache
} // namespace tfk
