// Fast path for the /infer request body.
//
// on the served path the time goes to JSON, not to the model -- nlohmann builds a DOM node for
// each of the 150,528 numbers of a full image (1-3 MB of text) and the handler then converts that tree into a
// vector<float> (src/worker_node.cpp:53-54).  This scanner reads the one shape every client sends,
//     { "request_id": "<plain ASCII>", "input_data": [ <JSON numbers> ] }      (either key order, any whitespace)
// straight into the vector, with the conversions nlohmann would apply: an integer literal goes through int64, any
// other number through a correctly rounded double, then static_cast<float>.
// Anything else -- escapes or non-ASCII in strings, other keys, duplicate keys, non-numeric elements, malformed
// numbers, 19+ digit integers, trailing bytes -- makes it return false WITHOUT judging the request: the caller then
// runs the ordinary nlohmann path, so acceptance, results and error texts are exactly what they were.
#pragma once
#include <charconv>
#include <cstring>
#include <string>
#include <vector>

namespace fast_request {

inline const char* skip_ws(const char* p, const char* e) {
  while (p < e && (*p == ' ' || *p == '\n' || *p == '\r' || *p == '\t')) ++p;
  return p;
}

// A plain string: printable ASCII without '"' or '\\'.  Returns the position after the closing quote or nullptr.
inline const char* plain_string(const char* p, const char* e, std::string& out) {
  if (p >= e || *p != '"') return nullptr;
  const char* s = ++p;
  while (p < e && *p != '"') {
    const unsigned char c = static_cast<unsigned char>(*p);
    if (c < 0x20 || c >= 0x80 || c == '\\') return nullptr;
    ++p;
  }
  if (p >= e) return nullptr;
  out.assign(s, size_t(p - s));
  return p + 1;
}

// One JSON number (RFC 8259 grammar) converted the way nlohmann + get<float>() would.
inline const char* number(const char* p, const char* e, float& out) {
  const char* s = p;
  if (p < e && *p == '-') ++p;
  if (p >= e) return nullptr;
  if (*p == '0') ++p;
  else if (*p >= '1' && *p <= '9') { while (p < e && *p >= '0' && *p <= '9') ++p; }
  else return nullptr;
  bool integral = true;
  if (p < e && *p == '.') {
    integral = false;
    ++p;
    if (p >= e || *p < '0' || *p > '9') return nullptr;
    while (p < e && *p >= '0' && *p <= '9') ++p;
  }
  if (p < e && (*p == 'e' || *p == 'E')) {
    integral = false;
    ++p;
    if (p < e && (*p == '+' || *p == '-')) ++p;
    if (p >= e || *p < '0' || *p > '9') return nullptr;
    while (p < e && *p >= '0' && *p <= '9') ++p;
  }
  if (integral) {
    if (p - s > 18) return nullptr;               // may not fit int64: let nlohmann pick uint64 / double
    long long v = 0;
    if (std::from_chars(s, p, v).ec != std::errc()) return nullptr;
    out = static_cast<float>(v);
  } else {
    double v = 0;
    auto r = std::from_chars(s, p, v);
    if (r.ec == std::errc::result_out_of_range) return nullptr;   // nlohmann raises out_of_range.406 for these
    if (r.ec != std::errc() || r.ptr != p) return nullptr;
    out = static_cast<float>(v);
  }
  return p;
}

inline bool parse(const std::string& body, std::string& request_id, std::vector<float>& input) {
  const char* p = body.data();
  const char* e = p + body.size();
  bool have_id = false, have_data = false;
  p = skip_ws(p, e);
  if (p >= e || *p != '{') return false;
  p = skip_ws(p + 1, e);
  std::string key;
  while (true) {
    p = plain_string(p, e, key);
    if (!p) return false;
    p = skip_ws(p, e);
    if (p >= e || *p != ':') return false;
    p = skip_ws(p + 1, e);
    if (key == "request_id" && !have_id) {
      p = plain_string(p, e, request_id);
      if (!p) return false;
      have_id = true;
    } else if (key == "input_data" && !have_data) {
      if (p >= e || *p != '[') return false;
      p = skip_ws(p + 1, e);
      input.clear();
      input.reserve(size_t(e - p) / 6 + 1);
      if (p < e && *p == ']') {
        ++p;
      } else {
        while (true) {
          float v;
          p = number(p, e, v);
          if (!p) return false;
          input.push_back(v);
          p = skip_ws(p, e);
          if (p >= e) return false;
          if (*p == ',') { p = skip_ws(p + 1, e); continue; }
          if (*p == ']') { ++p; break; }
          return false;
        }
      }
      have_data = true;
    } else {
      return false;
    }
    p = skip_ws(p, e);
    if (p >= e) return false;
    if (*p == ',') { p = skip_ws(p + 1, e); continue; }
    if (*p == '}') { ++p; break; }
    return false;
  }
  p = skip_ws(p, e);
  return p == e && have_id && have_data;
}

}  // namespace fast_request
