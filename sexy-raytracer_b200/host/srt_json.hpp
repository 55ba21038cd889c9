// srt_json.hpp — a small JSON reader for the glTF loader (objects, arrays, strings with escapes, numbers, literals).
#ifndef SRT_JSON_HPP_
#define SRT_JSON_HPP_

#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <string>
#include <vector>

namespace srt_json {

struct Value {
  enum Kind { Null, Bool, Number, String, Array, Object } kind = Null;
  bool b = false;
  double num = 0;
  std::string str;
  std::vector<Value> arr;
  std::vector<std::pair<std::string, Value>> obj;      // insertion order kept (cgltf walks attributes in file order)

  const Value* get(const char* key) const {
    if (kind != Object) return nullptr;
    for (const auto& kv : obj) if (kv.first == key) return &kv.second;
    return nullptr;
  }
  bool has(const char* key) const { return get(key) != nullptr; }
  double number(const char* key, double dflt) const { const Value* v = get(key); return v && v->kind == Number ? v->num : dflt; }
  long index(const char* key) const { const Value* v = get(key); return v && v->kind == Number ? (long)v->num : -1; }
  std::string string(const char* key) const { const Value* v = get(key); return v && v->kind == String ? v->str : std::string(); }
  size_t size() const { return kind == Array ? arr.size() : 0; }
  const Value& at(size_t i) const { static const Value none; return kind == Array && i < arr.size() ? arr[i] : none; }
};

class Parser {
 public:
  Parser(const char* s, size_t n) : p(s), e(s + n) {}
  bool parse(Value& out, std::string& err) {
    if (!value(out, 0)) { err = "JSON syntax error near byte " + std::to_string((size_t)(p - start())); return false; }
    ws();
    return true;
  }

 private:
  const char* p;
  const char* e;
  const char* s0 = nullptr;
  const char* start() { return s0 ? s0 : p; }
  void ws() { while (p < e && (*p == ' ' || *p == '\t' || *p == '\n' || *p == '\r')) ++p; }
  bool lit(const char* w) { size_t n = strlen(w); if ((size_t)(e - p) >= n && !memcmp(p, w, n)) { p += n; return true; } return false; }
  static void utf8(std::string& o, unsigned c) {
    if (c < 0x80) o += (char)c;
    else if (c < 0x800) { o += (char)(0xc0 | (c >> 6)); o += (char)(0x80 | (c & 0x3f)); }
    else if (c < 0x10000) { o += (char)(0xe0 | (c >> 12)); o += (char)(0x80 | ((c >> 6) & 0x3f)); o += (char)(0x80 | (c & 0x3f)); }
    else { o += (char)(0xf0 | (c >> 18)); o += (char)(0x80 | ((c >> 12) & 0x3f)); o += (char)(0x80 | ((c >> 6) & 0x3f)); o += (char)(0x80 | (c & 0x3f)); }
  }
  bool hex4(unsigned& c) {
    if (e - p < 4) return false;
    c = 0;
    for (int i = 0; i < 4; i++) {
      const char h = *p++;
      c <<= 4;
      if (h >= '0' && h <= '9') c |= h - '0'; else if (h >= 'a' && h <= 'f') c |= h - 'a' + 10;
      else if (h >= 'A' && h <= 'F') c |= h - 'A' + 10; else return false;
    }
    return true;
  }
  bool string(std::string& o) {
    if (p >= e || *p != '"') return false;
    ++p;
    while (p < e && *p != '"') {
      if (*p == '\\') {
        if (++p >= e) return false;
        const char c = *p++;
        switch (c) {
          case '"': o += '"'; break; case '\\': o += '\\'; break; case '/': o += '/'; break;
          case 'b': o += '\b'; break; case 'f': o += '\f'; break; case 'n': o += '\n'; break;
          case 'r': o += '\r'; break; case 't': o += '\t'; break;
          case 'u': {
            unsigned u;
            if (!hex4(u)) return false;
            if (u >= 0xd800 && u < 0xdc00 && e - p >= 6 && p[0] == '\\' && p[1] == 'u') {
              p += 2;
              unsigned lo;
              if (!hex4(lo)) return false;
              u = 0x10000 + ((u - 0xd800) << 10) + (lo - 0xdc00);
            }
            utf8(o, u);
            break;
          }
          default: return false;
        }
      } else {
        o += *p++;
      }
    }
    if (p >= e) return false;
    ++p;
    return true;
  }
  bool value(Value& v, int depth) {
    if (!s0) s0 = p;
    if (depth > 64) return false;
    ws();
    if (p >= e) return false;
    if (*p == '{') {
      ++p; v.kind = Value::Object; ws();
      if (p < e && *p == '}') { ++p; return true; }
      for (;;) {
        ws();
        std::string k;
        if (!string(k)) return false;
        ws();
        if (p >= e || *p++ != ':') return false;
        v.obj.emplace_back(std::move(k), Value());
        if (!value(v.obj.back().second, depth + 1)) return false;
        ws();
        if (p < e && *p == ',') { ++p; continue; }
        if (p < e && *p == '}') { ++p; return true; }
        return false;
      }
    }
    if (*p == '[') {
      ++p; v.kind = Value::Array; ws();
      if (p < e && *p == ']') { ++p; return true; }
      for (;;) {
        v.arr.emplace_back();
        if (!value(v.arr.back(), depth + 1)) return false;
        ws();
        if (p < e && *p == ',') { ++p; continue; }
        if (p < e && *p == ']') { ++p; return true; }
        return false;
      }
    }
    if (*p == '"') { v.kind = Value::String; return string(v.str); }
    if (lit("true")) { v.kind = Value::Bool; v.b = true; return true; }
    if (lit("false")) { v.kind = Value::Bool; v.b = false; return true; }
    if (lit("null")) { v.kind = Value::Null; return true; }
    // number: copy the token so strtod cannot run past the buffer
    const char* q = p;
    while (q < e && (strchr("+-.eE", *q) || (*q >= '0' && *q <= '9'))) ++q;
    if (q == p) return false;
    std::string tok(p, q);
    char* endp = nullptr;
    v.num = strtod(tok.c_str(), &endp);
    if (endp != tok.c_str() + tok.size()) return false;
    v.kind = Value::Number;
    p = q;
    return true;
  }
};

}  // namespace srt_json
#endif
