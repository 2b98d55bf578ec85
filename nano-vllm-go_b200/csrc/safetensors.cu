// safetensors.cu — nvllm_load_safetensors: checkpoint files -> device weights, directly.
//
// Replaces tensor.LoadModelFromDirectory's weight half (purego/tensor/generic_loader.go:184-265
// single file, :1016-1163 sharded index; per-block :268-604) for the four families the
// reference maps (:60-182).  The reference decodes every tensor to fp32 on the host (a 7 B
// model becomes 28 GB of Go slices), transposes PyTorch [out, in] linears to [in, out]
// (Transpose, :362-456,531-545), de-interleaves fused QKV (:674-765) and concatenates
// [gate | up] (:531-536).  Here the files are mmap'ed, each tensor is copied to the device in
// its file dtype (F32 / F16 / BF16 stay narrow, :645-663 lists the same three) and every one of
// those transforms is index arithmetic of the device-side repack (weights.cu): a PyTorch linear
// already has the device orientation, so it is a strided row copy; GPT-2's Conv1D [in, out]
// tensors take the transposing path; Falcon's fused query_key_value rows are [Q heads | K | V],
// which IS the fused device matrix.
//
// Lookup rules follow the reference: a missing name is retried with a "transformer." prefix
// (:621-629); norm biases are "<key with .weight -> .bias>" (:601); the FFN biases are looked
// up as "<FFNUpKey>.bias" = "...c_fc.weight.bias", which no checkpoint has, so GPT-2's MLP
// biases are never loaded (:559-560 — reproduced on purpose, SURVEY.md §8a "quirks"); a
// missing lm_head falls back to the tied embedding (:255-259).
#include <fcntl.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <map>
#include <memory>
#include <string>
#include <vector>

#include "common.cuh"
#include "handle.h"

namespace nvllm {
namespace {

// ---------------------------------------------------------------- minimal JSON
// Enough for a safetensors header and model.safetensors.index.json: objects, arrays, strings
// (with escapes), numbers, true / false / null.
struct Json {
  enum Kind { NUL, BOOL, NUM, STR, ARR, OBJ } kind = NUL;
  double num = 0;
  bool b = false;
  std::string str;
  std::vector<Json> arr;
  std::vector<std::pair<std::string, Json>> obj;
  const Json* get(const std::string& k) const {
    for (auto& kv : obj)
      if (kv.first == k) return &kv.second;
    return nullptr;
  }
};

struct JsonParser {
  const char* p;
  const char* end;
  std::string err;
  void ws() {
    while (p < end && (*p == ' ' || *p == '\n' || *p == '\t' || *p == '\r')) p++;
  }
  bool fail(const char* m) {
    if (err.empty()) err = m;
    return false;
  }
  bool string(std::string* out) {
    if (p >= end || *p != '"') return fail("expected string");
    p++;
    out->clear();
    while (p < end && *p != '"') {
      if (*p == '\\') {
        if (++p >= end) return fail("bad escape");
        switch (*p) {
          case 'n': out->push_back('\n'); break;
          case 't': out->push_back('\t'); break;
          case 'r': out->push_back('\r'); break;
          case 'b': out->push_back('\b'); break;
          case 'f': out->push_back('\f'); break;
          case 'u': {
            if (end - p < 5) return fail("bad \\u escape");
            unsigned cp = 0;
            for (int i = 1; i <= 4; i++) {
              const char c = p[i];
              cp = cp * 16 + (c >= '0' && c <= '9' ? c - '0' : (c | 32) >= 'a' && (c | 32) <= 'f' ? (c | 32) - 'a' + 10 : 0);
            }
            p += 4;
            if (cp < 0x80) out->push_back((char)cp);
            else if (cp < 0x800) { out->push_back((char)(0xC0 | cp >> 6)); out->push_back((char)(0x80 | (cp & 63))); }
            else { out->push_back((char)(0xE0 | cp >> 12)); out->push_back((char)(0x80 | ((cp >> 6) & 63))); out->push_back((char)(0x80 | (cp & 63))); }
            break;
          }
          default: out->push_back(*p);
        }
        p++;
      } else {
        out->push_back(*p++);
      }
    }
    if (p >= end) return fail("unterminated string");
    p++;
    return true;
  }
  bool value(Json* v, int depth = 0) {
    if (depth > 64) return fail("nesting too deep");
    ws();
    if (p >= end) return fail("unexpected end");
    if (*p == '{') {
      v->kind = Json::OBJ;
      p++;
      ws();
      if (p < end && *p == '}') { p++; return true; }
      for (;;) {
        ws();
        std::string k;
        if (!string(&k)) return false;
        ws();
        if (p >= end || *p != ':') return fail("expected ':'");
        p++;
        v->obj.emplace_back(std::move(k), Json());
        if (!value(&v->obj.back().second, depth + 1)) return false;
        ws();
        if (p < end && *p == ',') { p++; continue; }
        if (p < end && *p == '}') { p++; return true; }
        return fail("expected ',' or '}'");
      }
    }
    if (*p == '[') {
      v->kind = Json::ARR;
      p++;
      ws();
      if (p < end && *p == ']') { p++; return true; }
      for (;;) {
        v->arr.emplace_back();
        if (!value(&v->arr.back(), depth + 1)) return false;
        ws();
        if (p < end && *p == ',') { p++; continue; }
        if (p < end && *p == ']') { p++; return true; }
        return fail("expected ',' or ']'");
      }
    }
    if (*p == '"') {
      v->kind = Json::STR;
      return string(&v->str);
    }
    if (end - p >= 4 && !strncmp(p, "true", 4)) { v->kind = Json::BOOL; v->b = true; p += 4; return true; }
    if (end - p >= 5 && !strncmp(p, "false", 5)) { v->kind = Json::BOOL; v->b = false; p += 5; return true; }
    if (end - p >= 4 && !strncmp(p, "null", 4)) { v->kind = Json::NUL; p += 4; return true; }
    char* e = nullptr;
    std::string tmp(p, std::min<size_t>(end - p, 48));
    v->num = strtod(tmp.c_str(), &e);
    if (e == tmp.c_str()) return fail("unexpected character");
    v->kind = Json::NUM;
    p += e - tmp.c_str();
    return true;
  }
};

// ---------------------------------------------------------------- files
struct MappedFile {
  void* base = MAP_FAILED;
  size_t size = 0;
  ~MappedFile() {
    if (base != MAP_FAILED) munmap(base, size);
  }
  bool open_ro(const std::string& path, std::string* err) {
    const int fd = open(path.c_str(), O_RDONLY);
    if (fd < 0) { *err = "cannot open " + path; return false; }
    struct stat st;
    if (fstat(fd, &st) != 0 || st.st_size < 8) { close(fd); *err = path + ": too short for a safetensors file"; return false; }
    size = (size_t)st.st_size;
    base = mmap(nullptr, size, PROT_READ, MAP_PRIVATE, fd, 0);
    close(fd);
    if (base == MAP_FAILED) { *err = "mmap failed for " + path; return false; }
    madvise(base, size, MADV_SEQUENTIAL);
    return true;
  }
};

struct TensorRef {
  int dtype = -1;  // nvllm_dtype
  std::vector<int64_t> shape;
  const uint8_t* data = nullptr;
  uint64_t nbytes = 0;
  uint64_t numel() const {
    uint64_t n = 1;
    for (int64_t d : shape) n *= (uint64_t)d;
    return n;
  }
};

struct Checkpoint {
  std::vector<std::unique_ptr<MappedFile>> files;
  std::map<std::string, TensorRef> tensors;

  // 8-byte little-endian header length, JSON header, then the data section (offsets are
  // relative to it): generic_loader.go:192-199
  bool add_file(const std::string& path, std::string* err) {
    std::unique_ptr<MappedFile> f(new MappedFile());
    if (!f->open_ro(path, err)) return false;
    const uint8_t* b = (const uint8_t*)f->base;
    uint64_t hlen = 0;
    for (int i = 7; i >= 0; i--) hlen = (hlen << 8) | b[i];
    if (hlen > f->size - 8) { *err = path + ": header length exceeds the file"; return false; }
    JsonParser jp{(const char*)b + 8, (const char*)b + 8 + hlen, ""};
    Json root;
    if (!jp.value(&root) || root.kind != Json::OBJ) { *err = path + ": header is not a JSON object (" + jp.err + ")"; return false; }
    const uint8_t* data = b + 8 + hlen;
    const uint64_t dsize = f->size - 8 - hlen;
    for (auto& kv : root.obj) {
      if (kv.first == "__metadata__") continue;
      const Json* dt = kv.second.get("dtype");
      const Json* sh = kv.second.get("shape");
      const Json* off = kv.second.get("data_offsets");
      if (!dt || !sh || !off || dt->kind != Json::STR || sh->kind != Json::ARR || off->kind != Json::ARR || off->arr.size() != 2) {
        *err = path + ": malformed entry for tensor " + kv.first;
        return false;
      }
      TensorRef t;
      t.dtype = dt->str == "F32" ? NVLLM_DT_F32 : dt->str == "BF16" ? NVLLM_DT_BF16 : dt->str == "F16" ? NVLLM_DT_F16 : -1;
      for (auto& d : sh->arr) t.shape.push_back((int64_t)d.num);
      const uint64_t a = (uint64_t)off->arr[0].num, e = (uint64_t)off->arr[1].num;
      if (a > e || e > dsize) { *err = path + ": data_offsets of " + kv.first + " outside the file"; return false; }
      t.data = data + a;
      t.nbytes = e - a;
      if (t.dtype >= 0 && t.nbytes != t.numel() * (t.dtype == NVLLM_DT_F32 ? 4 : 2)) {
        *err = path + ": " + kv.first + " has " + std::to_string(t.nbytes) + " bytes for its shape";
        return false;
      }
      tensors[kv.first] = std::move(t);  // unsupported dtypes stay listed: the error names them on use
    }
    files.push_back(std::move(f));
    return true;
  }

  // model.safetensors.index.json -> the shard files of weight_map (generic_loader.go:1042-1105),
  // else model.safetensors (:1037-1039)
  bool open_dir(const std::string& dir, std::string* err) {
    const std::string idx = dir + "/model.safetensors.index.json";
    struct stat st;
    if (stat(idx.c_str(), &st) == 0) {
      MappedFile f;
      if (!f.open_ro(idx, err)) return false;
      JsonParser jp{(const char*)f.base, (const char*)f.base + f.size, ""};
      Json root;
      if (!jp.value(&root) || root.kind != Json::OBJ) { *err = idx + ": not a JSON object (" + jp.err + ")"; return false; }
      const Json* wm = root.get("weight_map");
      if (!wm || wm->kind != Json::OBJ) { *err = idx + ": no weight_map"; return false; }
      std::vector<std::string> shards;
      for (auto& kv : wm->obj) {
        if (kv.second.kind != Json::STR) continue;
        bool seen = false;
        for (auto& s : shards) seen = seen || s == kv.second.str;
        if (!seen) shards.push_back(kv.second.str);
      }
      for (auto& s : shards)
        if (!add_file(dir + "/" + s, err)) return false;
      return true;
    }
    return add_file(dir + "/model.safetensors", err);
  }

  // generic_loader.go:621-629: the name, then "transformer." + name
  const TensorRef* find(const std::string& name) const {
    auto it = tensors.find(name);
    if (it == tensors.end()) it = tensors.find("transformer." + name);
    return it == tensors.end() ? nullptr : &it->second;
  }
};

// generic_loader.go:60-182, the fields this path needs
struct Mapping {
  const char *tok, *pos, *lm_head, *prefix, *q, *out, *ffn_up, *ffn_down, *router, *moe_in, *moe_out,
      *attn_norm, *ffn_norm, *input_norm, *final_norm;
  bool qkv_combined, pytorch;
};
const Mapping kGPT2 = {"wte.weight", "wpe.weight", "", "h.%d", ".attn.c_attn.weight", ".attn.c_proj.weight",
                       ".mlp.c_fc.weight", ".mlp.c_proj.weight", "", "", "", ".ln_1.weight", ".ln_2.weight", "",
                       "ln_f.weight", true, false};
const Mapping kFalcon = {"transformer.word_embeddings.weight", "", "lm_head.weight", "transformer.h.%d",
                         ".self_attention.query_key_value.weight", ".self_attention.dense.weight",
                         ".mlp.dense_h_to_4h.weight", ".mlp.dense_4h_to_h.weight", "", "", "", "", "",
                         ".input_layernorm.weight", "transformer.ln_f.weight", true, true};
const Mapping kLlama = {"model.embed_tokens.weight", "", "lm_head.weight", "model.layers.%d",
                        ".self_attn.q_proj.weight", ".self_attn.o_proj.weight", ".mlp.gate_proj.weight",
                        ".mlp.down_proj.weight", "", "", "", ".input_layernorm.weight",
                        ".post_attention_layernorm.weight", "", "model.norm.weight", false, true};
const Mapping kGraniteMoE = {"model.embed_tokens.weight", "", "", "model.layers.%d", ".self_attn.q_proj.weight",
                             ".self_attn.o_proj.weight", "", "", ".block_sparse_moe.router.layer.weight",
                             ".block_sparse_moe.input_linear.weight", ".block_sparse_moe.output_linear.weight",
                             ".input_layernorm.weight", ".post_attention_layernorm.weight", "",
                             "model.norm.weight", false, true};

std::string replace_once(std::string s, const std::string& from, const std::string& to) {
  const size_t i = s.find(from);
  if (i != std::string::npos) s.replace(i, from.size(), to);
  return s;
}

struct Loader {
  nvllm_handle* h;
  const Checkpoint& ck;
  uint64_t bytes = 0;
  int tensors = 0;

  int put(int id, int layer, const TensorRef& t, const WeightSrc& src, const std::string& name,
          const uint8_t* data = nullptr, uint64_t numel = 0) {
    if (t.dtype < 0) return set_err(h, NVLLM_ERR_INVALID, "tensor %s has an unsupported dtype (F32 / F16 / BF16 only)", name.c_str());
    const uint64_t n = numel ? numel : t.numel();
    int rc = set_weight_src(h, id, layer, t.dtype, data ? data : t.data, n, src);
    if (rc) {
      std::string e = h->err;
      return set_err(h, rc, "%s: %s", name.c_str(), e.c_str());
    }
    bytes += n * (t.dtype == NVLLM_DT_F32 ? 4 : 2);
    tensors++;
    return NVLLM_OK;
  }
  int required(int id, int layer, const std::string& name, const WeightSrc& src) {
    const TensorRef* t = ck.find(name);
    if (!t) return set_err(h, NVLLM_ERR_INVALID, "required tensor '%s' not found (also tried transformer.%s)", name.c_str(), name.c_str());
    return put(id, layer, *t, src, name);
  }
  int optional(int id, int layer, const std::string& name, const WeightSrc& src) {
    const TensorRef* t = ck.find(name);
    return t ? put(id, layer, *t, src, name) : NVLLM_OK;
  }
  // loadNorm (:595-604): weight required, "<key: .weight -> .bias>" optional
  int norm(int id_w, int id_b, int layer, const std::string& key) {
    int rc = required(id_w, layer, key, WeightSrc());
    if (rc) return rc;
    return optional(id_b, layer, replace_once(key, ".weight", ".bias"), WeightSrc());
  }
};

}  // namespace

int load_safetensors(nvllm_handle* h, const char* dir, int arch, uint64_t* bytes_out, int* tensors_out) {
  if (h->finalized) return set_err(h, NVLLM_ERR_STATE, "load_safetensors after finalize");
  const Mapping* mp = nullptr;
  switch (arch) {
    case NVLLM_ARCH_GPT2: mp = &kGPT2; break;
    case NVLLM_ARCH_FALCON: mp = &kFalcon; break;
    case NVLLM_ARCH_LLAMA: mp = &kLlama; break;
    case NVLLM_ARCH_GRANITE_MOE: mp = &kGraniteMoE; break;
    default: return set_err(h, NVLLM_ERR_INVALID, "unsupported architecture %d", arch);
  }
  const Mapping& m = *mp;
  const nvllm_config& c = h->cfg;
  const bool mqa = c.attention_type == NVLLM_ATTN_MQA;
  if (arch == NVLLM_ARCH_FALCON && !mqa)
    return set_err(h, NVLLM_ERR_INVALID, "the Falcon mapping needs a multi-query config");
  if ((arch == NVLLM_ARCH_GRANITE_MOE) != (c.use_moe != 0))
    return set_err(h, NVLLM_ERR_INVALID, "mixture-of-experts config and mapping disagree");
  Checkpoint ck;
  std::string err;
  if (!ck.open_dir(dir, &err)) return set_err(h, NVLLM_ERR_INVALID, "%s", err.c_str());
  Loader ld{h, ck};
  WeightSrc lin;  // a linear as this family's files hold it
  lin.out_in = m.pytorch;
  const int H = c.hidden, hd = c.head_dim, nh = c.num_heads;
  int rc;
#define TRY(expr) \
  if ((rc = (expr)) != NVLLM_OK) return rc;

  TRY(ld.required(NVLLM_W_TOKEN_EMBEDDING, -1, m.tok, WeightSrc()));
  if (m.pos[0] && c.position_type == NVLLM_POS_LEARNED) TRY(ld.required(NVLLM_W_POS_EMBEDDING, -1, m.pos, WeightSrc()));
  for (int l = 0; l < c.num_layers; l++) {
    char pre[128];
    snprintf(pre, sizeof pre, m.prefix, l);
    const std::string p(pre);
    // ---- attention (loadAttention, :353-456)
    if (m.qkv_combined && mqa) {
      // Falcon query_key_value [(n_h + 2) hd, H]: rows [Q heads | K | V] (splitFalconQKV :705-748
      // de-interleaves the transposed copy; in file orientation the split is two row ranges)
      const std::string name = p + m.q;
      const TensorRef* t = ck.find(name);
      if (!t) return set_err(h, NVLLM_ERR_INVALID, "required tensor '%s' not found", name.c_str());
      const uint64_t qn = (uint64_t)nh * hd * H, kvn = (uint64_t)2 * hd * H;
      if (t->numel() != qn + kvn)
        return set_err(h, NVLLM_ERR_INVALID, "%s: %llu elements, expected %llu", name.c_str(),
                       (unsigned long long)t->numel(), (unsigned long long)(qn + kvn));
      const size_t esz = t->dtype == NVLLM_DT_F32 ? 4 : 2;
      TRY(ld.put(NVLLM_W_Q, l, *t, lin, name, t->data, qn));
      TRY(ld.put(NVLLM_W_KV, l, *t, lin, name, t->data + qn * esz, kvn));
    } else if (m.qkv_combined) {
      // GPT-2 c_attn [H, 3H], Conv1D orientation [in, out]: Q | K | V are column ranges
      // (splitGPT2QKV :674-702); c_attn.bias [3H] likewise (:420-428)
      const std::string name = p + m.q;
      const TensorRef* t = ck.find(name);
      if (!t) return set_err(h, NVLLM_ERR_INVALID, "required tensor '%s' not found", name.c_str());
      const int ids[3] = {NVLLM_W_Q, NVLLM_W_K, NVLLM_W_V};
      const int bids[3] = {NVLLM_W_Q_BIAS, NVLLM_W_K_BIAS, NVLLM_W_V_BIAS};
      const TensorRef* tb = ck.find(replace_once(name, ".weight", ".bias"));
      for (int i = 0; i < 3; i++) {
        WeightSrc s;
        s.ld = (size_t)3 * H;
        s.col0 = (size_t)i * H;
        TRY(ld.put(ids[i], l, *t, s, name));
        if (tb) TRY(ld.put(bids[i], l, *tb, s, name + " (bias)"));
      }
    } else {
      TRY(ld.required(NVLLM_W_Q, l, p + m.q, lin));
      TRY(ld.required(NVLLM_W_K, l, p + ".self_attn.k_proj.weight", lin));
      TRY(ld.required(NVLLM_W_V, l, p + ".self_attn.v_proj.weight", lin));
    }
    TRY(ld.required(NVLLM_W_O, l, p + m.out, lin));
    if (c.attention_type == NVLLM_ATTN_MHA)  // :434-436, MHA only
      TRY(ld.optional(NVLLM_W_O_BIAS, l, p + replace_once(m.out, ".weight", ".bias"), WeightSrc()));
    // ---- FFN / MoE (loadFFN :513-563, loadMoE :566-592)
    if (c.use_moe) {
      WeightSrc r;
      r.out_in = true;  // router [E, H] is always transposed to [H, E] by the reference (:575)
      TRY(ld.required(NVLLM_W_MOE_ROUTER, l, p + m.router, r));
      TRY(ld.required(NVLLM_W_MOE_INPUT, l, p + m.moe_in, WeightSrc()));
      TRY(ld.required(NVLLM_W_MOE_OUTPUT, l, p + m.moe_out, WeightSrc()));
    } else {
      const std::string up = p + m.ffn_up;
      bool done = false;
      if (c.activation_type == NVLLM_ACT_SWIGLU) {
        const std::string upk = replace_once(up, "gate_proj", "up_proj");
        const TensorRef *g = ck.find(up), *u = ck.find(upk);
        if (g && u && upk != up) {  // separate gate / up (:523-536)
          TRY(ld.put(W_INT_FFN_GATE, l, *g, lin, up));
          TRY(ld.put(W_INT_FFN_UP, l, *u, lin, upk));
          done = true;
        }
      }
      if (!done) TRY(ld.required(NVLLM_W_FFN_W1, l, up, lin));
      TRY(ld.required(NVLLM_W_FFN_W2, l, p + m.ffn_down, lin));
      // :559-560 — the keys the reference builds end in ".weight.bias"
      if (c.activation_type != NVLLM_ACT_SWIGLU) TRY(ld.optional(NVLLM_W_FFN_B1, l, up + ".bias", WeightSrc()));
      TRY(ld.optional(NVLLM_W_FFN_B2, l, p + m.ffn_down + ".bias", WeightSrc()));
    }
    // ---- norms (:325-349)
    if (c.block_style == NVLLM_BLOCK_PARALLEL) {
      TRY(ld.norm(NVLLM_W_ATTN_NORM_W, NVLLM_W_ATTN_NORM_B, l, p + m.input_norm));
    } else {
      TRY(ld.norm(NVLLM_W_ATTN_NORM_W, NVLLM_W_ATTN_NORM_B, l, p + (m.attn_norm[0] ? m.attn_norm : m.input_norm)));
      TRY(ld.norm(NVLLM_W_FFN_NORM_W, NVLLM_W_FFN_NORM_B, l, p + m.ffn_norm));
    }
  }
  TRY(ld.norm(NVLLM_W_FINAL_NORM_W, NVLLM_W_FINAL_NORM_B, -1, m.final_norm));
  if (!c.tied_embedding) {
    // :242-259 — lm_head when the mapping names one and the file has it, else the embedding
    const TensorRef* t = m.lm_head[0] ? ck.find(m.lm_head) : nullptr;
    WeightSrc s;
    s.out_in = true;  // [V, H]: PyTorch orientation; the embedding matrix has it by construction
    if (t) {
      if (!m.pytorch) s.out_in = false;
      TRY(ld.put(NVLLM_W_LM_HEAD, -1, *t, s, m.lm_head));
    } else {
      const TensorRef* e = ck.find(m.tok);
      TRY(ld.put(NVLLM_W_LM_HEAD, -1, *e, s, std::string(m.tok) + " (as LM head)"));
    }
  }
#undef TRY
  if (bytes_out) *bytes_out = ld.bytes;
  if (tensors_out) *tensors_out = ld.tensors;
  return NVLLM_OK;
}

}  // namespace nvllm

using namespace nvllm;

extern "C" int nvllm_load_safetensors(nvllm_handle* h, const char* model_dir, int arch,
                                      uint64_t* bytes_read, int32_t* tensors_loaded) {
  if (!h) return NVLLM_ERR_INVALID;
  if (!model_dir) return set_err(h, NVLLM_ERR_INVALID, "null model directory");
  NV_CHECK(h, cudaSetDevice(h->dev));
  int n = 0;
  int rc = load_safetensors(h, model_dir, arch, bytes_read, &n);
  if (tensors_loaded) *tensors_loaded = n;
  return rc;
}

// Host-only: look a tensor up the way the loader does (index / single file, "transformer."
// retry).  Returns NVLLM_OK and fills the outputs, NVLLM_ERR_INVALID when the directory cannot
// be read or the tensor is absent.  No device, no handle: usable to validate a checkpoint.
extern "C" int nvllm_safetensors_find(const char* model_dir, const char* name, int32_t* dtype,
                                      int32_t* ndim, int64_t shape[8], uint64_t* nbytes,
                                      uint64_t* first_bytes) {
  if (!model_dir || !name) return NVLLM_ERR_INVALID;
  Checkpoint ck;
  std::string err;
  if (!ck.open_dir(model_dir, &err)) {
    g_create_error = err;
    return NVLLM_ERR_INVALID;
  }
  const TensorRef* t = ck.find(name);
  if (!t) {
    g_create_error = std::string("tensor not found: ") + name;
    return NVLLM_ERR_INVALID;
  }
  if (dtype) *dtype = t->dtype;
  if (ndim) *ndim = (int32_t)t->shape.size();
  if (shape)
    for (size_t i = 0; i < 8; i++) shape[i] = i < t->shape.size() ? t->shape[i] : 0;
  if (nbytes) *nbytes = t->nbytes;
  if (first_bytes) {
    *first_bytes = 0;
    memcpy(first_bytes, t->data, std::min<uint64_t>(8, t->nbytes));
  }
  return NVLLM_OK;
}
