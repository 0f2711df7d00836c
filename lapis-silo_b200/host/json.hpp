// Minimal JSON value / parser / writer for the host side of silo_b200.
// The reference uses nlohmann::json (query.cpp:13-28, query_result.cpp:10-25); this is a
// from-scratch, dependency-free equivalent that keeps the number-kind distinction the
// reference's validation relies on (is_number_unsigned / is_number_integer / is_number_float).
#pragma once
#include <charconv>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace silo::json {

struct ParseError : std::runtime_error {
   using std::runtime_error::runtime_error;
};

class Value {
  public:
   enum Kind { NUL, BOOL, UINT, INT, FLOAT, STRING, ARRAY, OBJECT };
   Kind kind = NUL;
   bool b = false;
   uint64_t u = 0;
   int64_t i = 0;
   double d = 0;
   std::string s;
   std::vector<Value> arr;
   // insertion-ordered object; lookups are linear (queries are tiny)
   std::vector<std::pair<std::string, Value>> obj;

   Value() = default;
   static Value null() { return Value(); }
   static Value boolean(bool v) { Value x; x.kind = BOOL; x.b = v; return x; }
   static Value integer(int64_t v) {
      Value x;
      if (v >= 0) { x.kind = UINT; x.u = static_cast<uint64_t>(v); x.i = v; }
      else { x.kind = INT; x.i = v; }
      x.d = static_cast<double>(v);
      return x;
   }
   static Value number(double v) { Value x; x.kind = FLOAT; x.d = v; return x; }
   static Value string(std::string v) { Value x; x.kind = STRING; x.s = std::move(v); return x; }
   static Value array() { Value x; x.kind = ARRAY; return x; }
   static Value object() { Value x; x.kind = OBJECT; return x; }

   bool is_null() const { return kind == NUL; }
   bool is_boolean() const { return kind == BOOL; }
   bool is_string() const { return kind == STRING; }
   bool is_array() const { return kind == ARRAY; }
   bool is_object() const { return kind == OBJECT; }
   bool is_number() const { return kind == UINT || kind == INT || kind == FLOAT; }
   bool is_number_unsigned() const { return kind == UINT; }
   bool is_number_integer() const { return kind == UINT || kind == INT; }
   bool is_number_float() const { return kind == FLOAT; }

   bool contains(const std::string& key) const {
      if (kind != OBJECT) return false;
      for (const auto& kv : obj) if (kv.first == key) return true;
      return false;
   }
   const Value& at(const std::string& key) const {
      for (const auto& kv : obj) if (kv.first == key) return kv.second;
      throw ParseError("key not found: " + key);
   }
   const Value& operator[](const std::string& key) const { return at(key); }
   void set(const std::string& key, Value v) {
      for (auto& kv : obj) if (kv.first == key) { kv.second = std::move(v); return; }
      obj.emplace_back(key, std::move(v));
   }
   double as_double() const {
      if (kind == FLOAT) return d;
      if (kind == UINT) return static_cast<double>(u);
      if (kind == INT) return static_cast<double>(i);
      throw ParseError("not a number");
   }
   int64_t as_int() const {
      if (kind == UINT) return static_cast<int64_t>(u);
      if (kind == INT) return i;
      if (kind == FLOAT) return static_cast<int64_t>(d);
      throw ParseError("not a number");
   }
   const std::string& as_string() const {
      if (kind != STRING) throw ParseError("not a string");
      return s;
   }
   std::string dump() const { std::string out; write(out); return out; }

   void write(std::string& out) const {
      switch (kind) {
         case NUL: out += "null"; break;
         case BOOL: out += b ? "true" : "false"; break;
         case UINT: out += std::to_string(u); break;
         case INT: out += std::to_string(i); break;
         case FLOAT: writeDouble(out, d); break;
         case STRING: writeString(out, s); break;
         case ARRAY: {
            out += '[';
            for (size_t k = 0; k < arr.size(); ++k) { if (k) out += ','; arr[k].write(out); }
            out += ']';
            break;
         }
         case OBJECT: {
            out += '{';
            for (size_t k = 0; k < obj.size(); ++k) {
               if (k) out += ',';
               writeString(out, obj[k].first);
               out += ':';
               obj[k].second.write(out);
            }
            out += '}';
            break;
         }
      }
   }
   // Shortest form that parses back to the same double, written like printf("%.{n}g"): n = the number of significant
   // digits of the shortest round-trip representation (std::to_chars) but at least 15, i.e. the first of %.15g, %.16g, %.17g that round-trips.
   static void writeDouble(std::string& out, double d) {
      if (std::isnan(d) || std::isinf(d)) { out += "null"; return; }
      char tmp[48], buf[48];
      const auto shortest = std::to_chars(tmp, tmp + sizeof tmp, d, std::chars_format::scientific);
      int digits = 0;
      for (const char* c = tmp; c < shortest.ptr && *c != 'e'; ++c) digits += *c >= '0' && *c <= '9';
      const auto r = std::to_chars(buf, buf + sizeof buf, d, std::chars_format::general, std::max(digits, 15));
      out.append(buf, r.ptr);
      bool integral = true;
      for (const char* c = buf; c < r.ptr; ++c) if (*c == '.' || *c == 'e' || *c == 'E') { integral = false; break; }
      if (integral) out += ".0";
   }
   static void writeString(std::string& out, const std::string& str) {
      out += '"';
      for (unsigned char c : str) {
         switch (c) {
            case '"': out += "\\\""; break;
            case '\\': out += "\\\\"; break;
            case '\n': out += "\\n"; break;
            case '\r': out += "\\r"; break;
            case '\t': out += "\\t"; break;
            default:
               if (c < 0x20) { char buf[8]; snprintf(buf, sizeof buf, "\\u%04x", c); out += buf; }
               else out += static_cast<char>(c);
         }
      }
      out += '"';
   }
};

class Parser {
   const char* p;
   const char* end;
   [[noreturn]] void fail(const std::string& what) const { throw ParseError("JSON parse error: " + what); }
   void ws() { while (p < end && (*p == ' ' || *p == '\n' || *p == '\t' || *p == '\r')) ++p; }
   static void appendUtf8(std::string& out, uint32_t cp) {
      if (cp < 0x80) out += static_cast<char>(cp);
      else if (cp < 0x800) { out += static_cast<char>(0xC0 | (cp >> 6)); out += static_cast<char>(0x80 | (cp & 0x3F)); }
      else if (cp < 0x10000) { out += static_cast<char>(0xE0 | (cp >> 12)); out += static_cast<char>(0x80 | ((cp >> 6) & 0x3F)); out += static_cast<char>(0x80 | (cp & 0x3F)); }
      else { out += static_cast<char>(0xF0 | (cp >> 18)); out += static_cast<char>(0x80 | ((cp >> 12) & 0x3F)); out += static_cast<char>(0x80 | ((cp >> 6) & 0x3F)); out += static_cast<char>(0x80 | (cp & 0x3F)); }
   }
   uint32_t hex4() {
      if (end - p < 4) fail("bad \\u escape");
      uint32_t v = 0;
      for (int k = 0; k < 4; ++k) {
         char c = *p++;
         v <<= 4;
         if (c >= '0' && c <= '9') v |= c - '0';
         else if (c >= 'a' && c <= 'f') v |= c - 'a' + 10;
         else if (c >= 'A' && c <= 'F') v |= c - 'A' + 10;
         else fail("bad hex digit");
      }
      return v;
   }
   std::string parseString() {
      ++p;  // opening quote
      std::string out;
      while (true) {
         if (p >= end) fail("unterminated string");
         char c = *p++;
         if (c == '"') break;
         if (c == '\\') {
            if (p >= end) fail("bad escape");
            char e = *p++;
            switch (e) {
               case '"': out += '"'; break;
               case '\\': out += '\\'; break;
               case '/': out += '/'; break;
               case 'b': out += '\b'; break;
               case 'f': out += '\f'; break;
               case 'n': out += '\n'; break;
               case 'r': out += '\r'; break;
               case 't': out += '\t'; break;
               case 'u': {
                  uint32_t cp = hex4();
                  if (cp >= 0xD800 && cp < 0xDC00 && end - p >= 6 && p[0] == '\\' && p[1] == 'u') {
                     p += 2;
                     uint32_t lo = hex4();
                     cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
                  }
                  appendUtf8(out, cp);
                  break;
               }
               default: fail("bad escape");
            }
         } else out += c;
      }
      return out;
   }
   Value parseNumber() {
      const char* start = p;
      bool neg = false, is_float = false;
      if (*p == '-') { neg = true; ++p; }
      while (p < end && ((*p >= '0' && *p <= '9') || *p == '.' || *p == 'e' || *p == 'E' || *p == '+' || *p == '-')) {
         if (*p == '.' || *p == 'e' || *p == 'E') is_float = true;
         ++p;
      }
      std::string text(start, p);
      if (text.empty() || text == "-") fail("bad number");
      Value v;
      if (is_float) { v.kind = Value::FLOAT; v.d = strtod(text.c_str(), nullptr); return v; }
      if (neg) {
         v.kind = Value::INT; v.i = strtoll(text.c_str(), nullptr, 10); v.d = static_cast<double>(v.i);
         if (v.i == 0) { v.kind = Value::UINT; v.u = 0; }  // "-0" parses as integer 0
      } else {
         v.kind = Value::UINT; v.u = strtoull(text.c_str(), nullptr, 10); v.i = static_cast<int64_t>(v.u); v.d = static_cast<double>(v.u);
      }
      return v;
   }
   Value parseValue(int depth) {
      if (depth > 512) fail("nesting too deep");
      ws();
      if (p >= end) fail("unexpected end");
      char c = *p;
      if (c == '{') {
         ++p;
         Value v = Value::object();
         ws();
         if (p < end && *p == '}') { ++p; return v; }
         while (true) {
            ws();
            if (p >= end || *p != '"') fail("expected key");
            std::string key = parseString();
            ws();
            if (p >= end || *p != ':') fail("expected ':'");
            ++p;
            v.set(key, parseValue(depth + 1));
            ws();
            if (p < end && *p == ',') { ++p; continue; }
            if (p < end && *p == '}') { ++p; break; }
            fail("expected ',' or '}'");
         }
         return v;
      }
      if (c == '[') {
         ++p;
         Value v = Value::array();
         ws();
         if (p < end && *p == ']') { ++p; return v; }
         while (true) {
            v.arr.push_back(parseValue(depth + 1));
            ws();
            if (p < end && *p == ',') { ++p; continue; }
            if (p < end && *p == ']') { ++p; break; }
            fail("expected ',' or ']'");
         }
         return v;
      }
      if (c == '"') return Value::string(parseString());
      if (c == 't' && end - p >= 4 && !strncmp(p, "true", 4)) { p += 4; return Value::boolean(true); }
      if (c == 'f' && end - p >= 5 && !strncmp(p, "false", 5)) { p += 5; return Value::boolean(false); }
      if (c == 'n' && end - p >= 4 && !strncmp(p, "null", 4)) { p += 4; return Value(); }
      if (c == '-' || (c >= '0' && c <= '9')) return parseNumber();
      fail(std::string("unexpected character '") + c + "'");
   }

  public:
   Parser(const char* begin, const char* end_) : p(begin), end(end_) {}
   Value parseDocument() {
      Value v = parseValue(0);
      ws();
      if (p != end) fail("trailing characters");
      return v;
   }
};

inline Value parse(const std::string& text) { return Parser(text.data(), text.data() + text.size()).parseDocument(); }
inline Value parse(const char* begin, size_t len) { return Parser(begin, begin + len).parseDocument(); }

}  // namespace silo::json
