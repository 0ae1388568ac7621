// Shard: append-only key/value record file `<folder>/shard.dat` plus a batch reader that decodes image records
// straight into caller-provided (pinned) batch buffers.
//
// Record layout (byte-compatible with the reference on x86-64: include/utils/shard.h:33-142,
// src/utils/shard.cc:24-206 — the length fields are size_t, i.e. 8 bytes, not the uint32 its header comment claims):
//     [u64 keylen][key bytes][u64 vallen][value bytes] ...
// Modes: 'r' read, 'c' create (truncate), 'a' append.  Opening for append re-scans the file, logically truncates
// it after the last COMPLETE tuple (a crashed writer may have left a torn tail) and rebuilds the key set so
// duplicate keys are ignored — the loader is restartable.
//
// This is host-only code; it lives in libsinga_b200.so so the data path has no Python per-record cost.
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <sys/stat.h>
#include <sys/types.h>
#include <unistd.h>

#include <string>
#include <unordered_set>
#include <vector>

#include "../include/singa_b200.h"
#include "common.h"

namespace sg {

struct Shard {
  std::string path;
  char mode = 'r';
  FILE* f = nullptr;
  std::unordered_set<std::string> keys;
  std::vector<char> iobuf;
  std::string key, val;   // last tuple read
  long long count_cache = -1;
};

static bool read_u64(FILE* f, uint64_t* v) { return fread(v, sizeof(uint64_t), 1, f) == 1; }

// Reads one tuple; returns 1 ok, 0 clean EOF, -1 torn tuple.
static int read_tuple(FILE* f, std::string* key, std::string* val, bool want_val) {
  uint64_t klen = 0, vlen = 0;
  if (!read_u64(f, &klen)) return feof(f) ? 0 : -1;
  if (klen > (1ull << 20)) return -1;
  key->resize(klen);
  if (klen && fread(&(*key)[0], 1, klen, f) != klen) return -1;
  if (!read_u64(f, &vlen)) return -1;
  if (vlen > (1ull << 34)) return -1;
  if (want_val) {
    val->resize(vlen);
    if (vlen && fread(&(*val)[0], 1, vlen, f) != vlen) return -1;
  } else {
    if (fseeko(f, static_cast<off_t>(vlen), SEEK_CUR) != 0) return -1;
    // fseeko past EOF succeeds: verify the tuple is really complete
    const off_t pos = ftello(f);
    struct stat st;
    if (fstat(fileno(f), &st) != 0 || pos > st.st_size) return -1;
  }
  return 1;
}

// ---- minimal protobuf wire decoding of singa.Record{ image = 2 { shape = 1*, label = 2, pixel = 3, data = 4* } }
static bool get_varint(const uint8_t*& p, const uint8_t* end, uint64_t* out) {
  uint64_t v = 0;
  int shift = 0;
  while (p < end && shift < 64) {
    const uint8_t b = *p++;
    v |= static_cast<uint64_t>(b & 0x7F) << shift;
    if (!(b & 0x80)) { *out = v; return true; }
    shift += 7;
  }
  return false;
}
static bool skip_field(const uint8_t*& p, const uint8_t* end, int wire) {
  uint64_t t;
  switch (wire) {
    case 0: return get_varint(p, end, &t);
    case 1: if (end - p < 8) return false; p += 8; return true;
    case 2: if (!get_varint(p, end, &t) || static_cast<uint64_t>(end - p) < t) return false; p += t; return true;
    case 5: if (end - p < 4) return false; p += 4; return true;
    default: return false;
  }
}
// Destination of one decoded image: a uint8 row or an fp32 row of `dim` values.  Byte-valued records widen losslessly
// into an fp32 row; FLOAT-valued records (mean-subtracted images: negative and fractional values,
// src/worker/layer.cc:595-598 copies them unchanged) only fit an fp32 row — asked to narrow them into bytes the
// decoder records `saw_float` and the batch reader fails instead of clamping.
struct PixelSink {
  uint8_t* u8;
  float* f32;
  long long dim;
  bool saw_float;
  void put_bytes(const uint8_t* src, long long n) {
    const long long c = n < dim ? n : dim;
    if (u8 != nullptr) {
      memcpy(u8, src, c);
      if (c < dim) memset(u8 + c, 0, dim - c);
    } else {
      for (long long i = 0; i < c; ++i) f32[i] = static_cast<float>(src[i]);
      for (long long i = c; i < dim; ++i) f32[i] = 0.f;
    }
  }
  void put_float(long long idx, float v) {
    if (f32 != nullptr) { if (idx < dim) f32[idx] = v; }
    else saw_float = true;
  }
};

// Decodes one Record into `sink` and `label`.
static bool decode_image_record(const std::string& rec, PixelSink& sink, int* label) {
  const long long dim = sink.dim;
  (void)dim;
  const uint8_t* p = reinterpret_cast<const uint8_t*>(rec.data());
  const uint8_t* end = p + rec.size();
  bool got = false;
  while (p < end) {
    uint64_t tag;
    if (!get_varint(p, end, &tag)) return false;
    const int field = static_cast<int>(tag >> 3), wire = static_cast<int>(tag & 7);
    if (field == 2 && wire == 2) {
      uint64_t len;
      if (!get_varint(p, end, &len) || static_cast<uint64_t>(end - p) < len) return false;
      const uint8_t* q = p;
      const uint8_t* qend = p + len;
      p = qend;
      long long nfloat = 0;
      while (q < qend) {
        uint64_t t2;
        if (!get_varint(q, qend, &t2)) return false;
        const int f2 = static_cast<int>(t2 >> 3), w2 = static_cast<int>(t2 & 7);
        if (f2 == 2 && w2 == 0) {
          uint64_t v;
          if (!get_varint(q, qend, &v)) return false;
          *label = static_cast<int>(v);
        } else if (f2 == 3 && w2 == 2) {
          uint64_t n;
          if (!get_varint(q, qend, &n) || static_cast<uint64_t>(qend - q) < n) return false;
          sink.put_bytes(q, static_cast<long long>(n));
          q += n;
          got = true;
        } else if (f2 == 4 && (w2 == 5 || w2 == 2)) {
          uint64_t n = 4;
          if (w2 == 2 && (!get_varint(q, qend, &n) || static_cast<uint64_t>(qend - q) < n)) return false;
          for (uint64_t o = 0; o + 4 <= n; o += 4) {
            float v;
            memcpy(&v, q + o, 4);
            sink.put_float(nfloat++, v);
          }
          q += n;
          got = true;
        } else if (!skip_field(q, qend, w2)) {
          return false;
        }
      }
    } else if (!skip_field(p, end, wire)) {
      return false;
    }
  }
  return got;
}

}  // namespace sg

using namespace sg;

extern "C" void* sg_shard_open(const char* folder, char mode, long long capacity) {
  if (folder == nullptr || (mode != 'r' && mode != 'c' && mode != 'a')) {
    set_error("sg_shard_open: mode must be 'r', 'c' or 'a'");
    return nullptr;
  }
  Shard* s = new Shard;
  s->path = std::string(folder) + "/shard.dat";
  s->mode = mode;
  if (mode != 'r') mkdir(folder, 0755);
  if (mode == 'a') {
    // scan for the last complete tuple, remember the keys
    long long good = 0;
    FILE* r = fopen(s->path.c_str(), "rb");
    if (r != nullptr) {
      std::string k, v;
      while (true) {
        const int rc = read_tuple(r, &k, &v, false);
        if (rc != 1) break;
        s->keys.insert(k);
        good = ftello(r);
      }
      fclose(r);
      if (truncate(s->path.c_str(), good) != 0) { /* file may not exist yet */ }
    }
    s->f = fopen(s->path.c_str(), r != nullptr ? "r+b" : "wb");
    if (s->f != nullptr) fseeko(s->f, good, SEEK_SET);
  } else {
    s->f = fopen(s->path.c_str(), mode == 'r' ? "rb" : "wb");
  }
  if (s->f == nullptr) {
    set_error("sg_shard_open: cannot open %s", s->path.c_str());
    delete s;
    return nullptr;
  }
  if (capacity < (1 << 12)) capacity = 1 << 12;
  if (capacity > (256ll << 20)) capacity = 256ll << 20;
  s->iobuf.resize(static_cast<size_t>(capacity));
  setvbuf(s->f, s->iobuf.data(), _IOFBF, s->iobuf.size());
  return s;
}

extern "C" void sg_shard_close(void* h) {
  Shard* s = static_cast<Shard*>(h);
  if (s == nullptr) return;
  if (s->f) fclose(s->f);
  delete s;
}

// 1 = inserted, 0 = duplicate key ignored, <0 = error
extern "C" int sg_shard_insert(void* h, const void* key, long long klen, const void* val, long long vlen) {
  Shard* s = static_cast<Shard*>(h);
  if (s == nullptr || s->mode == 'r' || klen <= 0 || vlen < 0) { set_error("sg_shard_insert: bad arguments"); return -1; }
  std::string k(static_cast<const char*>(key), klen);
  if (!s->keys.insert(k).second) return 0;
  const uint64_t kl = klen, vl = vlen;
  if (fwrite(&kl, 8, 1, s->f) != 1 || fwrite(key, 1, klen, s->f) != static_cast<size_t>(klen) ||
      fwrite(&vl, 8, 1, s->f) != 1 || (vlen && fwrite(val, 1, vlen, s->f) != static_cast<size_t>(vlen))) {
    set_error("sg_shard_insert: write failed");
    return -2;
  }
  s->count_cache = -1;
  return 1;
}

extern "C" int sg_shard_flush(void* h) {
  Shard* s = static_cast<Shard*>(h);
  if (s == nullptr || s->f == nullptr) return -1;
  return fflush(s->f) == 0 ? 0 : -1;
}

// 1 = a tuple is available through sg_shard_key/sg_shard_val, 0 = end of file
extern "C" int sg_shard_next(void* h) {
  Shard* s = static_cast<Shard*>(h);
  if (s == nullptr || s->mode != 'r') { set_error("sg_shard_next: shard not opened for reading"); return -1; }
  return read_tuple(s->f, &s->key, &s->val, true) == 1 ? 1 : 0;
}
extern "C" const void* sg_shard_key(void* h, long long* len) {
  Shard* s = static_cast<Shard*>(h);
  *len = static_cast<long long>(s->key.size());
  return s->key.data();
}
extern "C" const void* sg_shard_val(void* h, long long* len) {
  Shard* s = static_cast<Shard*>(h);
  *len = static_cast<long long>(s->val.size());
  return s->val.data();
}
extern "C" void sg_shard_seek_to_first(void* h) {
  Shard* s = static_cast<Shard*>(h);
  if (s && s->f) { clearerr(s->f); fseeko(s->f, 0, SEEK_SET); }
}
extern "C" long long sg_shard_count(void* h) {
  Shard* s = static_cast<Shard*>(h);
  if (s == nullptr) return -1;
  if (s->mode != 'r') return static_cast<long long>(s->keys.size());
  if (s->count_cache >= 0) return s->count_cache;
  FILE* r = fopen(s->path.c_str(), "rb");
  if (r == nullptr) return -1;
  long long n = 0;
  std::string k, v;
  while (read_tuple(r, &k, &v, false) == 1) ++n;
  fclose(r);
  s->count_cache = n;
  return n;
}

// Fill pixels[batch][dim] (uint8) and labels[batch] from the next `batch` image records, wrapping around at
// the end of the shard (ShardDataLayer semantics, src/worker/layer.cc:646-661).  Returns records read or <0.
static int shard_read_batch(void* h, uint8_t* px8, float* px32, int* labels, int batch, long long dim) {
  Shard* s = static_cast<Shard*>(h);
  if (s == nullptr || s->mode != 'r') { set_error("sg_shard_read_image_batch: shard not opened for reading"); return -1; }
  for (int i = 0; i < batch; ++i) {
    int rc = read_tuple(s->f, &s->key, &s->val, true);
    if (rc != 1) {
      sg_shard_seek_to_first(h);
      rc = read_tuple(s->f, &s->key, &s->val, true);
      if (rc != 1) { set_error("sg_shard_read_image_batch: empty shard"); return -2; }
    }
    int label = 0;
    PixelSink sink{px8 ? px8 + static_cast<long long>(i) * dim : nullptr, px32 ? px32 + static_cast<long long>(i) * dim : nullptr,
                   dim, false};
    if (!decode_image_record(s->val, sink, &label)) {
      set_error("sg_shard_read_image_batch: record %s is not a SingleLabelImageRecord", s->key.c_str());
      return -3;
    }
    if (sink.saw_float) {
      set_error("sg_shard_read_image_batch: record %s holds float pixels (e.g. a mean-subtracted image); they do not fit a "
                "uint8 batch - read it with sg_shard_read_image_batch_f32", s->key.c_str());
      return -4;
    }
    labels[i] = label;
  }
  return batch;
}

extern "C" int sg_shard_read_image_batch(void* h, void* pixels, int* labels, int batch, long long dim) {
  return shard_read_batch(h, static_cast<uint8_t*>(pixels), nullptr, labels, batch, dim);
}

extern "C" int sg_shard_read_image_batch_f32(void* h, float* pixels, int* labels, int batch, long long dim) {
  return shard_read_batch(h, nullptr, pixels, labels, batch, dim);
}
