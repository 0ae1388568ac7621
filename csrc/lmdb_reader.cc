// Read-only LMDB environment reader (no liblmdb on the box): walks the B+tree of `<dir>/data.mdb` through an mmap
// and decodes Caffe `Datum` values straight into caller-provided (pinned) batch buffers — the input side of the
// reference's kLMDBData layer (src/worker/layer.cc:237-328: mdb_env_open / mdb_cursor_get(MDB_NEXT), wrap around at
// the end, Datum -> SingleLabelImageRecord).
//
// On-disk format (LMDB 0.9, 64-bit little endian), restated from the format itself:
//   page header (16 B): u64 pgno | u16 pad | u16 flags | { u16 lower, u16 upper }  or  { u32 overflow_pages }
//       flags: 0x01 branch, 0x02 leaf, 0x04 overflow, 0x08 meta
//       u16 node offsets follow the header; number of nodes = (lower - 16) / 2
//   meta pages 0 and 1 (body at +16): u32 magic 0xBEEFC0DE | u32 version | u64 address | u64 mapsize |
//       2 x db record (48 B: u32 pad(= page size in the FREE db) | u16 flags | u16 depth | u64 branch_pages |
//       u64 leaf_pages | u64 overflow_pages | u64 entries | u64 root) | u64 last_pg | u64 txnid
//       the meta with the larger txnid is current; db 1 is the main database
//   node (8 B header): u16 lo | u16 hi | u16 flags | u16 ksize | key | data
//       leaf: data size = lo | hi << 16; flags & 0x01 => data is a u64 page number of an overflow run whose bytes
//       start 16 B into the first page;  branch: child page = lo | hi << 16 | flags << 32
#include <fcntl.h>
#include <stdint.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <string>
#include <vector>

#include "../include/singa_b200.h"
#include "common.h"

namespace sg {

constexpr uint32_t LMDB_MAGIC = 0xBEEFC0DEu;
constexpr int PAGEHDR = 16;
constexpr uint16_t P_BRANCH = 0x01, P_LEAF = 0x02, P_OVERFLOW = 0x04, P_META = 0x08;
constexpr uint16_t F_BIGDATA = 0x01;
constexpr uint64_t P_INVALID = ~0ull;

struct LmdbEnv {
  int fd = -1;
  const uint8_t* map = nullptr;
  size_t map_len = 0;
  uint32_t psize = 4096;
  uint64_t root = P_INVALID;
  uint64_t entries = 0;
  uint64_t last_pg = 0;
  int depth = 0;
  // cursor: one (page, index) per tree level
  std::vector<uint64_t> pg;
  std::vector<int> idx;
  bool positioned = false;
  const uint8_t* key = nullptr; uint64_t klen = 0;
  const uint8_t* val = nullptr; uint64_t vlen = 0;
};

template <typename T> static inline T rd(const uint8_t* p) { T v; memcpy(&v, p, sizeof(T)); return v; }

static const uint8_t* page_ptr(const LmdbEnv* e, uint64_t pgno) {
  if (pgno > e->last_pg || (pgno + 1) * static_cast<uint64_t>(e->psize) > e->map_len) return nullptr;
  return e->map + pgno * static_cast<uint64_t>(e->psize);
}
static inline int page_nkeys(const uint8_t* pg) { return (static_cast<int>(rd<uint16_t>(pg + 12)) - PAGEHDR) / 2; }
static inline uint16_t page_flags(const uint8_t* pg) { return rd<uint16_t>(pg + 10); }
static const uint8_t* node_ptr(const LmdbEnv* e, const uint8_t* pg, int i) {
  const uint16_t off = rd<uint16_t>(pg + PAGEHDR + 2 * i);
  if (off < PAGEHDR || off + 8u > e->psize) return nullptr;
  return pg + off;
}

// Loads the node under the cursor (deepest level must be a leaf).
static bool load_current(LmdbEnv* e) {
  const uint8_t* pg = page_ptr(e, e->pg.back());
  if (pg == nullptr || !(page_flags(pg) & P_LEAF)) return false;
  const uint8_t* n = node_ptr(e, pg, e->idx.back());
  if (n == nullptr) return false;
  const uint32_t dsize = rd<uint16_t>(n) | (static_cast<uint32_t>(rd<uint16_t>(n + 2)) << 16);
  const uint16_t flags = rd<uint16_t>(n + 4);
  const uint16_t ksize = rd<uint16_t>(n + 6);
  const uint8_t* page_end = pg + e->psize;
  if (n + 8 + ksize > page_end) return false;
  e->key = n + 8; e->klen = ksize;
  if (flags & F_BIGDATA) {
    if (n + 8 + ksize + 8 > page_end) return false;
    const uint64_t opg = rd<uint64_t>(n + 8 + ksize);
    const uint8_t* ov = page_ptr(e, opg);
    if (ov == nullptr || !(page_flags(ov) & P_OVERFLOW)) return false;
    const uint64_t npages = rd<uint32_t>(ov + 12);
    if (npages == 0 || static_cast<uint64_t>(dsize) + PAGEHDR > npages * e->psize) return false;
    if ((opg + npages) * static_cast<uint64_t>(e->psize) > e->map_len) return false;
    e->val = ov + PAGEHDR; e->vlen = dsize;
  } else {
    if (n + 8 + ksize + dsize > page_end) return false;
    e->val = n + 8 + ksize; e->vlen = dsize;
  }
  return true;
}

// Descends from level `lvl` (whose page/index are set) always taking the first child, down to a leaf.
static bool descend_first(LmdbEnv* e, size_t lvl) {
  e->pg.resize(lvl + 1); e->idx.resize(lvl + 1);
  for (int guard = 0; guard < 64; ++guard) {
    const uint8_t* pg = page_ptr(e, e->pg[lvl]);
    if (pg == nullptr) return false;
    const uint16_t fl = page_flags(pg);
    if (fl & P_LEAF) return page_nkeys(pg) > 0 && e->idx[lvl] < page_nkeys(pg);
    if (!(fl & P_BRANCH) || e->idx[lvl] >= page_nkeys(pg)) return false;
    const uint8_t* n = node_ptr(e, pg, e->idx[lvl]);
    if (n == nullptr) return false;
    const uint64_t child = rd<uint16_t>(n) | (static_cast<uint64_t>(rd<uint16_t>(n + 2)) << 16) |
                           (static_cast<uint64_t>(rd<uint16_t>(n + 4)) << 32);
    e->pg.push_back(child); e->idx.push_back(0);
    ++lvl;
  }
  return false;
}

static bool cursor_first(LmdbEnv* e) {
  e->positioned = false;
  if (e->root == P_INVALID || e->entries == 0) return false;
  e->pg.assign(1, e->root); e->idx.assign(1, 0);
  if (!descend_first(e, 0) || !load_current(e)) return false;
  e->positioned = true;
  return true;
}

// 1 = moved, 0 = end of database, -1 = corrupt
static int cursor_next(LmdbEnv* e) {
  if (!e->positioned) return cursor_first(e) ? 1 : (e->entries == 0 ? 0 : -1);
  int lvl = static_cast<int>(e->pg.size()) - 1;
  while (lvl >= 0) {
    const uint8_t* pg = page_ptr(e, e->pg[lvl]);
    if (pg == nullptr) return -1;
    if (e->idx[lvl] + 1 < page_nkeys(pg)) {
      ++e->idx[lvl];
      if (!descend_first(e, static_cast<size_t>(lvl))) return -1;
      return load_current(e) ? 1 : -1;
    }
    --lvl;
  }
  e->positioned = false;
  return 0;
}

static bool read_meta(LmdbEnv* e, const uint8_t* pg, uint64_t* txnid, uint64_t* root, uint64_t* entries,
                      uint64_t* last_pg, int* depth, uint32_t* psize) {
  if (!(page_flags(pg) & P_META)) return false;
  const uint8_t* m = pg + PAGEHDR;
  if (rd<uint32_t>(m) != LMDB_MAGIC || rd<uint32_t>(m + 4) != 1) return false;
  const uint8_t* db0 = m + 24;          // FREE db: md_pad carries the page size
  const uint8_t* db1 = db0 + 48;        // MAIN db
  *psize = rd<uint32_t>(db0);
  *depth = rd<uint16_t>(db1 + 6);
  *entries = rd<uint64_t>(db1 + 32);
  *root = rd<uint64_t>(db1 + 40);
  *last_pg = rd<uint64_t>(db1 + 48);
  *txnid = rd<uint64_t>(db1 + 56);
  return true;
}

}  // namespace sg

using namespace sg;

namespace {

bool get_varint(const uint8_t*& p, const uint8_t* end, uint64_t* out) {
  uint64_t v = 0;
  for (int shift = 0; p < end && shift < 64; shift += 7) {
    const uint8_t b = *p++;
    v |= static_cast<uint64_t>(b & 0x7F) << shift;
    if (!(b & 0x80)) { *out = v; return true; }
  }
  return false;
}

// Caffe Datum { channels = 1, height = 2, width = 3, data = 4 (bytes), label = 5, float_data = 6, encoded = 7 }.
// pixels may be null (shape / label query only).
// `pixels` (uint8 row) or `pixels_f32` (fp32 row) receives the image; both may be null (shape / label query only).
// Float-valued Datums (float_data: mean-subtracted images, src/worker/layer.cc:289-293) only fit an fp32 row: asked to
// narrow them into bytes the decoder sets *saw_float and the batch reader fails instead of clamping.
bool decode_datum(const uint8_t* p, uint64_t len, uint8_t* pixels, long long dim, int* label, int* chw,
                  bool* encoded = nullptr, float* pixels_f32 = nullptr, bool* saw_float = nullptr) {
  const uint8_t* end = p + len;
  long long nfloat = 0;
  bool got = false;
  while (p < end) {
    uint64_t tag, v;
    if (!get_varint(p, end, &tag)) return false;
    const int field = static_cast<int>(tag >> 3), wire = static_cast<int>(tag & 7);
    if (wire == 0) {
      if (!get_varint(p, end, &v)) return false;
      if (field >= 1 && field <= 3 && chw != nullptr) chw[field - 1] = static_cast<int>(v);
      if (field == 5 && label != nullptr) *label = static_cast<int>(v);
      if (field == 7 && encoded != nullptr) *encoded = v != 0;
    } else if (wire == 2) {
      if (!get_varint(p, end, &v) || static_cast<uint64_t>(end - p) < v) return false;
      if (field == 4 && v > 0) {
        const long long c = static_cast<long long>(v) < dim ? static_cast<long long>(v) : dim;
        if (pixels != nullptr) {
          memcpy(pixels, p, c);
          if (c < dim) memset(pixels + c, 0, dim - c);
        } else if (pixels_f32 != nullptr) {
          for (long long i = 0; i < c; ++i) pixels_f32[i] = static_cast<float>(p[i]);
          for (long long i = c; i < dim; ++i) pixels_f32[i] = 0.f;
        }
        got = true;
      } else if (field == 6) {          // packed float_data
        for (uint64_t o = 0; o + 4 <= v; o += 4) {
          float f;
          memcpy(&f, p + o, 4);
          if (pixels_f32 != nullptr) { if (nfloat < dim) pixels_f32[nfloat] = f; }
          else if (saw_float != nullptr) *saw_float = true;
          ++nfloat;
        }
        got = got || v >= 4;
      }
      p += v;
    } else if (wire == 5) {
      if (end - p < 4) return false;
      if (field == 6) {                 // unpacked float_data element
        float f;
        memcpy(&f, p, 4);
        if (pixels_f32 != nullptr) { if (nfloat < dim) pixels_f32[nfloat] = f; }
        else if (saw_float != nullptr) *saw_float = true;
        ++nfloat;
        got = true;
      }
      p += 4;
    } else if (wire == 1) {
      if (end - p < 8) return false;
      p += 8;
    } else {
      return false;
    }
  }
  return got;
}

}  // namespace

extern "C" void* sg_lmdb_open(const char* dir) {
  std::string path = std::string(dir);
  struct stat st;
  if (stat(path.c_str(), &st) == 0 && S_ISDIR(st.st_mode)) path += "/data.mdb";
  const int fd = open(path.c_str(), O_RDONLY);
  if (fd < 0) { set_error("sg_lmdb_open: cannot open %s", path.c_str()); return nullptr; }
  if (fstat(fd, &st) != 0 || st.st_size < 2 * 512) {
    close(fd);
    set_error("sg_lmdb_open: %s is too small to be an LMDB data file", path.c_str());
    return nullptr;
  }
  void* m = mmap(nullptr, static_cast<size_t>(st.st_size), PROT_READ, MAP_SHARED, fd, 0);
  if (m == MAP_FAILED) { close(fd); set_error("sg_lmdb_open: mmap of %s failed", path.c_str()); return nullptr; }
  LmdbEnv* e = new LmdbEnv();
  e->fd = fd; e->map = static_cast<const uint8_t*>(m); e->map_len = static_cast<size_t>(st.st_size);
  uint64_t txn[2] = {0, 0}, root[2], ent[2], last[2];
  int depth[2]; uint32_t ps[2] = {0, 0};
  bool ok[2];
  e->last_pg = ~0ull >> 1;   // page_ptr bound while reading the metas
  ok[0] = read_meta(e, e->map, &txn[0], &root[0], &ent[0], &last[0], &depth[0], &ps[0]);
  ok[1] = false;
  if (ok[0] && ps[0] >= 512 && static_cast<size_t>(ps[0]) * 2 <= e->map_len)
    ok[1] = read_meta(e, e->map + ps[0], &txn[1], &root[1], &ent[1], &last[1], &depth[1], &ps[1]);
  if (!ok[0] && !ok[1]) {
    set_error("sg_lmdb_open: %s has no valid LMDB meta page (magic/version mismatch)", path.c_str());
    munmap(m, e->map_len); close(fd); delete e;
    return nullptr;
  }
  const int use = (ok[1] && (!ok[0] || txn[1] > txn[0])) ? 1 : 0;
  e->psize = ps[0] ? ps[0] : ps[1];
  e->root = root[use]; e->entries = ent[use]; e->last_pg = last[use]; e->depth = depth[use];
  return e;
}

extern "C" void sg_lmdb_close(void* h) {
  LmdbEnv* e = static_cast<LmdbEnv*>(h);
  if (e == nullptr) return;
  if (e->map != nullptr) munmap(const_cast<uint8_t*>(e->map), e->map_len);
  if (e->fd >= 0) close(e->fd);
  delete e;
}

extern "C" long long sg_lmdb_count(void* h) { return static_cast<long long>(static_cast<LmdbEnv*>(h)->entries); }

extern "C" void sg_lmdb_seek_to_first(void* h) { static_cast<LmdbEnv*>(h)->positioned = false; }

// 1 = a record is current, 0 = end of the database (the next call starts over), -1 = corrupt file
extern "C" int sg_lmdb_next(void* h) {
  const int rc = cursor_next(static_cast<LmdbEnv*>(h));
  if (rc < 0) set_error("sg_lmdb_next: corrupt page structure");
  return rc;
}

extern "C" const void* sg_lmdb_key(void* h, long long* len) {
  LmdbEnv* e = static_cast<LmdbEnv*>(h);
  *len = static_cast<long long>(e->klen);
  return e->key;
}

extern "C" const void* sg_lmdb_val(void* h, long long* len) {
  LmdbEnv* e = static_cast<LmdbEnv*>(h);
  *len = static_cast<long long>(e->vlen);
  return e->val;
}

// {channels, height, width} of the first record.
extern "C" int sg_lmdb_first_shape(void* h, int* chw) {
  LmdbEnv* e = static_cast<LmdbEnv*>(h);
  SG_REQUIRE(cursor_first(e), "sg_lmdb_first_shape: empty or corrupt database");
  chw[0] = chw[1] = chw[2] = 0;
  int label = 0;
  bool encoded = false;
  SG_REQUIRE(decode_datum(e->val, e->vlen, nullptr, 0, &label, chw, &encoded),
             "sg_lmdb_first_shape: first value is not a Datum");
  e->positioned = false;
  return encoded ? 2 : 0;
}

// Decodes the next `batch` Datum records (wrapping around at the end of the database, layer.cc:300-306) into
// pixels_u8[batch][dim] and labels[batch].
static int lmdb_read_batch(void* h, uint8_t* px8, float* px32, int* labels, int batch, long long dim) {
  LmdbEnv* e = static_cast<LmdbEnv*>(h);
  SG_REQUIRE(e->entries > 0, "sg_lmdb_read_datum_batch: empty database");
  for (int i = 0; i < batch; ++i) {
    int rc = cursor_next(e);
    if (rc == 0) rc = cursor_next(e);   // wrapped: start over
    SG_REQUIRE(rc == 1, "sg_lmdb_read_datum_batch: corrupt page structure");
    int label = 0;
    bool encoded = false, saw_float = false;
    SG_REQUIRE(decode_datum(e->val, e->vlen, px8 ? px8 + static_cast<long long>(i) * dim : nullptr, dim, &label, nullptr,
                            &encoded, px32 ? px32 + static_cast<long long>(i) * dim : nullptr, &saw_float),
               "sg_lmdb_read_datum_batch: record %d is not a Datum with pixel data", i);
    if (encoded) {
      // compressed image bytes (Datum.encoded): not decodable here; the caller re-reads this record through an image
      // decoder.  Step back so the record is not consumed.
      set_error("sg_lmdb_read_datum_batch: record %d holds an encoded image", i);
      return 2;
    }
    if (saw_float) {
      set_error("sg_lmdb_read_datum_batch: record %d holds float_data (e.g. a mean-subtracted image); it does not fit a "
                "uint8 batch - read it with sg_lmdb_read_datum_batch_f32", i);
      return 3;
    }
    labels[i] = label;
  }
  return 0;
}

// Decodes the next `batch` Datum records (wrapping around at the end of the database, layer.cc:300-306) into
// pixels[batch][dim] and labels[batch]; the _f32 form takes byte- and float-valued records alike.
extern "C" int sg_lmdb_read_datum_batch(void* h, void* pixels_u8, int* labels, int batch, long long dim) {
  return lmdb_read_batch(h, static_cast<uint8_t*>(pixels_u8), nullptr, labels, batch, dim);
}

extern "C" int sg_lmdb_read_datum_batch_f32(void* h, float* pixels, int* labels, int batch, long long dim) {
  return lmdb_read_batch(h, nullptr, pixels, labels, batch, dim);
}

// 1 when the first record carries float_data, 0 when bytes, <0 on error (lets a data layer size its batch buffers).
extern "C" int sg_lmdb_first_is_float(void* h) {
  LmdbEnv* e = static_cast<LmdbEnv*>(h);
  if (!cursor_first(e)) { set_error("sg_lmdb_first_is_float: empty or corrupt database"); return -1; }
  int label = 0;
  bool encoded = false, saw_float = false;
  uint8_t dummy = 0;
  const bool ok = decode_datum(e->val, e->vlen, &dummy, 0, &label, nullptr, &encoded, nullptr, &saw_float);
  e->positioned = false;
  if (!ok) { set_error("sg_lmdb_first_is_float: first value is not a Datum"); return -1; }
  return saw_float ? 1 : 0;
}
