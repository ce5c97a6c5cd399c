// update_stream.cpp -- see update_stream.hpp.
#include "update_stream.hpp"

#include <algorithm>
#include <atomic>
#include <charconv>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <functional>
#include <future>
#include <limits>
#include <sstream>
#include <thread>

namespace oryx {
namespace host {

namespace {

[[noreturn]] void bad(const std::string& why) { throw BadMessage(why); }

// ---------------------------------------------------------------- JSON (Jackson's default reader)
struct Cur {
  const char* p;
  const char* e;
  bool more() const { return p < e; }
};

inline void skipWs(Cur& c) {
  while (c.p < c.e && (*c.p == ' ' || *c.p == '\t' || *c.p == '\n' || *c.p == '\r')) c.p++;
}

void appendUtf8(std::string& out, uint32_t cp) {
  if (cp < 0x80) {
    out.push_back((char)cp);
  } else if (cp < 0x800) {
    out.push_back((char)(0xC0 | (cp >> 6)));
    out.push_back((char)(0x80 | (cp & 0x3F)));
  } else if (cp < 0x10000) {
    out.push_back((char)(0xE0 | (cp >> 12)));
    out.push_back((char)(0x80 | ((cp >> 6) & 0x3F)));
    out.push_back((char)(0x80 | (cp & 0x3F)));
  } else {
    out.push_back((char)(0xF0 | (cp >> 18)));
    out.push_back((char)(0x80 | ((cp >> 12) & 0x3F)));
    out.push_back((char)(0x80 | ((cp >> 6) & 0x3F)));
    out.push_back((char)(0x80 | (cp & 0x3F)));
  }
}

int hexVal(char ch) {
  if (ch >= '0' && ch <= '9') return ch - '0';
  if (ch >= 'a' && ch <= 'f') return ch - 'a' + 10;
  if (ch >= 'A' && ch <= 'F') return ch - 'A' + 10;
  return -1;
}

uint32_t readHex4(Cur& c) {
  if (c.e - c.p < 4) bad("truncated \\u escape");
  uint32_t v = 0;
  for (int i = 0; i < 4; i++) {
    int h = hexVal(c.p[i]);
    if (h < 0) bad("bad \\u escape");
    v = v * 16 + (uint32_t)h;
  }
  c.p += 4;
  return v;
}

// c.p at the opening quote; leaves c.p after the closing quote.
void readString(Cur& c, std::string& out) {
  out.clear();
  c.p++;
  for (;;) {
    const char* q = c.p;
    while (q < c.e && *q != '"' && *q != '\\' && (unsigned char)*q >= 0x20) q++;
    out.append(c.p, q);
    c.p = q;
    if (c.p >= c.e) bad("unterminated string");
    char ch = *c.p++;
    if (ch == '"') return;
    if ((unsigned char)ch < 0x20) bad("unescaped control character in string");
    if (c.p >= c.e) bad("truncated escape");
    char esc = *c.p++;
    switch (esc) {
      case '"': out.push_back('"'); break;
      case '\\': out.push_back('\\'); break;
      case '/': out.push_back('/'); break;
      case 'b': out.push_back('\b'); break;
      case 'f': out.push_back('\f'); break;
      case 'n': out.push_back('\n'); break;
      case 'r': out.push_back('\r'); break;
      case 't': out.push_back('\t'); break;
      case 'u': {
        uint32_t cp = readHex4(c);
        if (cp >= 0xD800 && cp < 0xDC00 && c.e - c.p >= 6 && c.p[0] == '\\' && c.p[1] == 'u') {
          Cur save = c;
          c.p += 2;
          uint32_t lo = readHex4(c);
          if (lo >= 0xDC00 && lo < 0xE000) cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
          else c = save;
        }
        appendUtf8(out, cp);
        break;
      }
      default: bad("unrecognized escape in string");
    }
  }
}

// JSON number grammar: -? (0 | [1-9][0-9]*) (\.[0-9]+)? ([eE][+-]?[0-9]+)?
struct NumTok {
  const char* b;
  const char* e;
  bool integral;
};
NumTok readNumberToken(Cur& c) {
  NumTok t{c.p, c.p, true};
  const char* p = c.p;
  if (p < c.e && *p == '-') p++;
  if (p >= c.e) bad("truncated number");
  if (*p == '0') {
    p++;
    if (p < c.e && *p >= '0' && *p <= '9') bad("leading zeros not allowed");
  } else if (*p >= '1' && *p <= '9') {
    while (p < c.e && *p >= '0' && *p <= '9') p++;
  } else {
    bad("unexpected character in numeric value");
  }
  if (p < c.e && *p == '.') {
    t.integral = false;
    p++;
    if (p >= c.e || *p < '0' || *p > '9') bad("decimal point not followed by a digit");
    while (p < c.e && *p >= '0' && *p <= '9') p++;
  }
  if (p < c.e && (*p == 'e' || *p == 'E')) {
    t.integral = false;
    p++;
    if (p < c.e && (*p == '+' || *p == '-')) p++;
    if (p >= c.e || *p < '0' || *p > '9') bad("exponent has no digits");
    while (p < c.e && *p >= '0' && *p <= '9') p++;
  }
  if (p < c.e && !(*p == ',' || *p == ']' || *p == '}' || *p == ' ' || *p == '\t' || *p == '\n' || *p == '\r'))
    bad("unexpected character after numeric value");
  t.e = p;
  c.p = p;
  return t;
}

// Double.parseDouble of a JSON number token: correctly rounded, overflow -> +-Infinity, underflow -> +-0.
double tokenToDouble(const NumTok& t) {
  double d = 0.0;
  auto r = std::from_chars(t.b, t.e, d);
  if (r.ec == std::errc() && r.ptr == t.e) return d;
  std::string tmp(t.b, t.e);
  return std::strtod(tmp.c_str(), nullptr);
}

// One pass over a JSON number: grammar check and value together.  Clinger's fast path -- at most 19
// significant digits collected into a 64-bit mantissa that fits 2^53, |decimal exponent| <= 22 -- gives
// the correctly rounded binary64 with one multiplication or division by an exact power of ten; anything
// else goes through from_chars on the token.  *integral tells "-0" (the Integer 0) from "-0.0".
double readNumber(Cur& c, bool* integral) {
  static const double kPow10[23] = {1e0,  1e1,  1e2,  1e3,  1e4,  1e5,  1e6,  1e7,  1e8,  1e9,  1e10, 1e11,
                                    1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};
  const char* const b = c.p;
  const char* p = c.p;
  const char* const e = c.e;
  bool neg = false;
  if (p < e && *p == '-') { neg = true; p++; }
  if (p >= e) bad("truncated number");
  uint64_t mant = 0;
  int digits = 0;       // significant digits taken into mant
  int dropped = 0;      // integer digits beyond the 19th (each scales by 10)
  bool exact = true;    // false once a non-zero digit had to be dropped
  int exp10 = 0;
  *integral = true;
  if (*p == '0') {
    p++;
    if (p < e && *p >= '0' && *p <= '9') bad("leading zeros not allowed");
  } else if (*p >= '1' && *p <= '9') {
    while (p < e && *p >= '0' && *p <= '9') {
      if (digits < 19) { mant = mant * 10 + (uint64_t)(*p - '0'); digits++; }
      else { dropped++; if (*p != '0') exact = false; }
      p++;
    }
  } else {
    bad("unexpected character in numeric value");
  }
  if (p < e && *p == '.') {
    *integral = false;
    p++;
    if (p >= e || *p < '0' || *p > '9') bad("decimal point not followed by a digit");
    while (p < e && *p >= '0' && *p <= '9') {
      if (digits < 19) {
        if (mant != 0 || *p != '0') { mant = mant * 10 + (uint64_t)(*p - '0'); digits++; }
        exp10--;
      } else if (*p != '0') {
        exact = false;
      }
      p++;
    }
  }
  exp10 += dropped;
  if (p < e && (*p == 'e' || *p == 'E')) {
    *integral = false;
    p++;
    bool eneg = false;
    if (p < e && (*p == '+' || *p == '-')) { eneg = *p == '-'; p++; }
    if (p >= e || *p < '0' || *p > '9') bad("exponent has no digits");
    int ev = 0;
    while (p < e && *p >= '0' && *p <= '9') {
      if (ev < 100000) ev = ev * 10 + (*p - '0');
      p++;
    }
    exp10 += eneg ? -ev : ev;
  }
  if (p < e && !(*p == ',' || *p == ']' || *p == '}' || *p == ' ' || *p == '\t' || *p == '\n' || *p == '\r'))
    bad("unexpected character after numeric value");
  c.p = p;
  if (exact && mant <= (1ull << 53) && exp10 >= -22 && exp10 <= 22) {
    double d = (double)mant;
    d = exp10 < 0 ? d / kPow10[-exp10] : d * kPow10[exp10];
    return neg ? -d : d;
  }
  NumTok t{b, p, *integral};
  return tokenToDouble(t);
}

bool matchLiteral(Cur& c, const char* lit) {
  size_t n = std::strlen(lit);
  if ((size_t)(c.e - c.p) < n || std::memcmp(c.p, lit, n) != 0) return false;
  c.p += n;
  return true;
}

void skipValue(Cur& c, int depth = 0) {
  if (depth > 512) bad("nesting too deep");
  skipWs(c);
  if (!c.more()) bad("unexpected end of input");
  char ch = *c.p;
  if (ch == '"') {
    std::string tmp;
    readString(c, tmp);
  } else if (ch == '[' || ch == '{') {
    const char close = ch == '[' ? ']' : '}';
    c.p++;
    skipWs(c);
    if (c.more() && *c.p == close) { c.p++; return; }
    for (;;) {
      if (ch == '{') {
        skipWs(c);
        if (!c.more() || *c.p != '"') bad("expected a field name");
        std::string tmp;
        readString(c, tmp);
        skipWs(c);
        if (!c.more() || *c.p != ':') bad("expected ':'");
        c.p++;
      }
      skipValue(c, depth + 1);
      skipWs(c);
      if (!c.more()) bad("unexpected end of input");
      if (*c.p == ',') { c.p++; continue; }
      if (*c.p == close) { c.p++; return; }
      bad("expected ',' or a closing bracket");
    }
  } else if (ch == '-' || (ch >= '0' && ch <= '9')) {
    readNumberToken(c);
  } else if (!(matchLiteral(c, "true") || matchLiteral(c, "false") || matchLiteral(c, "null"))) {
    bad("unrecognized token");
  }
}

// Object.toString() of what Jackson's untyped reader makes of a scalar: String, Integer/Long/BigInteger,
// Double, Boolean.
void readScalarAsString(Cur& c, std::string& out) {
  skipWs(c);
  if (!c.more()) bad("unexpected end of input");
  char ch = *c.p;
  if (ch == '"') {
    readString(c, out);
  } else if (ch == '-' || (ch >= '0' && ch <= '9')) {
    NumTok t = readNumberToken(c);
    if (t.integral) {
      out.assign(t.b, t.e);
      if (out == "-0") out = "0";
    } else {
      out = TextUtils::doubleToString(tokenToDouble(t));
    }
  } else if (matchLiteral(c, "true")) {
    out = "true";
  } else if (matchLiteral(c, "false")) {
    out = "false";
  } else if (matchLiteral(c, "null")) {
    bad("null where an id is required");  // NullPointerException on .toString()
  } else {
    bad("id must be a JSON scalar");
  }
}

// String element of a float[]: Jackson's _parseFloatPrimitive(String).
float stringToFloat(const std::string& raw) {
  size_t b = 0, e = raw.size();
  while (b < e && (unsigned char)raw[b] <= ' ') b++;
  while (e > b && (unsigned char)raw[e - 1] <= ' ') e--;
  std::string s = raw.substr(b, e - b);
  if (s.empty()) bad("empty string where a float is required");
  if (s == "NaN") return std::numeric_limits<float>::quiet_NaN();
  if (s == "Infinity" || s == "INF") return std::numeric_limits<float>::infinity();
  if (s == "-Infinity" || s == "-INF") return -std::numeric_limits<float>::infinity();
  const char* p = s.c_str();
  if (*p == '+') p++;
  float f = 0.0f;
  auto r = std::from_chars(p, s.c_str() + s.size(), f);
  if (r.ptr != s.c_str() + s.size()) bad("not a valid float value: " + s);
  if (r.ec != std::errc()) f = std::strtof(p, nullptr);
  return f;
}

// c.p at '['.  Appends the elements to `out` (number -> Double -> float, the narrowing convertValue does).
void readFloatArray(Cur& c, std::vector<float>& out) {
  skipWs(c);
  if (!c.more() || *c.p != '[') bad("expected an array of numbers");
  c.p++;
  skipWs(c);
  if (c.more() && *c.p == ']') { c.p++; return; }
  std::string tmp;
  for (;;) {
    skipWs(c);
    if (!c.more()) bad("unexpected end of input");
    char ch = *c.p;
    if (ch == '-' || (ch >= '0' && ch <= '9')) {
      bool integral;
      const double d = readNumber(c, &integral);
      out.push_back(integral && d == 0.0 ? 0.0f : (float)d);  // "-0" is the Integer 0, not -0.0
    } else if (ch == '"') {
      readString(c, tmp);
      out.push_back(stringToFloat(tmp));
    } else if (matchLiteral(c, "null")) {
      out.push_back(0.0f);
    } else {
      bad("array element is not a number");
    }
    skipWs(c);
    if (!c.more()) bad("unexpected end of input");
    if (*c.p == ',') { c.p++; continue; }
    if (*c.p == ']') { c.p++; return; }
    bad("expected ',' or ']'");
  }
}

void readStringArray(Cur& c, std::vector<std::string>& out, bool scalarsToString) {
  skipWs(c);
  if (!c.more() || *c.p != '[') bad("expected an array");
  c.p++;
  skipWs(c);
  if (c.more() && *c.p == ']') { c.p++; return; }
  for (;;) {
    skipWs(c);
    out.emplace_back();
    if (c.more() && *c.p == '"') {
      readString(c, out.back());
    } else if (scalarsToString) {
      if (c.more() && matchLiteral(c, "null")) bad("null element in a string array");
      readScalarAsString(c, out.back());
    } else {
      bad("array element is not a string");
    }
    skipWs(c);
    if (!c.more()) bad("unexpected end of input");
    if (*c.p == ',') { c.p++; continue; }
    if (*c.p == ']') { c.p++; return; }
    bad("expected ',' or ']'");
  }
}

// Shortest digits that round-trip, as (digits, exponent of the first digit).
template <typename T>
void shortestDigits(T v, std::string& digits, int& exp10) {
  char buf[64];
  auto r = std::to_chars(buf, buf + sizeof(buf), v, std::chars_format::scientific);
  std::string s(buf, r.ptr);  // d[.ddd]e[+-]xx
  size_t epos = s.find('e');
  exp10 = std::atoi(s.c_str() + epos + 1);
  digits.clear();
  for (size_t i = 0; i < epos; i++) if (s[i] != '.') digits.push_back(s[i]);
  if (digits.size() == 1) {
    // Float/Double.toString javadoc: when one digit suffices, the closest TWO-digit decimal that still
    // rounds to v is used (Float.MIN_VALUE prints as 1.4E-45, not 1.0E-45).
    r = std::to_chars(buf, buf + sizeof(buf), v, std::chars_format::scientific, 1);
    T back = 0;
    auto rb = std::from_chars(buf, r.ptr, back);
    if (rb.ec == std::errc() && back == v) {
      std::string s2(buf, r.ptr);
      size_t e2 = s2.find('e');
      exp10 = std::atoi(s2.c_str() + e2 + 1);
      digits.clear();
      for (size_t i = 0; i < e2; i++) if (s2[i] != '.') digits.push_back(s2[i]);
      while (digits.size() > 1 && digits.back() == '0') digits.pop_back();
    }
  }
}

// Layout rules of Float.toString / Double.toString: plain decimal for 1e-3 <= |v| < 1e7, else
// "computerized scientific notation"; always at least one digit either side of the point.
template <typename T>
std::string javaToString(T v) {
  if (std::isnan(v)) return "NaN";
  if (std::isinf(v)) return v > 0 ? "Infinity" : "-Infinity";
  std::string out;
  if (std::signbit(v)) { out.push_back('-'); v = -v; }
  if (v == 0) return out + "0.0";
  std::string d;
  int e;
  shortestDigits(v, d, e);
  if (e >= -3 && e < 7) {
    if (e >= 0) {
      for (int i = 0; i <= e; i++) out.push_back(i < (int)d.size() ? d[i] : '0');
      out.push_back('.');
      if ((int)d.size() > e + 1) out.append(d, e + 1, std::string::npos);
      else out.push_back('0');
    } else {
      out += "0.";
      out.append((size_t)(-e - 1), '0');
      out += d;
    }
  } else {
    out.push_back(d[0]);
    out.push_back('.');
    if (d.size() > 1) out.append(d, 1, std::string::npos);
    else out.push_back('0');
    out.push_back('E');
    out += std::to_string(e);
  }
  return out;
}

// Jackson's default string escaping on output.
void appendJSONString(std::string& out, const std::string& s) {
  static const char* hex = "0123456789ABCDEF";
  out.push_back('"');
  for (unsigned char ch : s) {
    switch (ch) {
      case '"': out += "\\\""; break;
      case '\\': out += "\\\\"; break;
      case '\b': out += "\\b"; break;
      case '\t': out += "\\t"; break;
      case '\n': out += "\\n"; break;
      case '\f': out += "\\f"; break;
      case '\r': out += "\\r"; break;
      default:
        if (ch < 0x20) {
          out += "\\u00";
          out.push_back(hex[ch >> 4]);
          out.push_back(hex[ch & 15]);
        } else {
          out.push_back((char)ch);
        }
    }
  }
  out.push_back('"');
}

void appendJSONFloats(std::string& out, const float* v, int n) {
  out.push_back('[');
  for (int i = 0; i < n; i++) {
    if (i) out.push_back(',');
    if (std::isnan(v[i]) || std::isinf(v[i])) {  // QUOTE_NON_NUMERIC_NUMBERS
      out.push_back('"');
      out += javaToString(v[i]);
      out.push_back('"');
    } else {
      out += javaToString(v[i]);
    }
  }
  out.push_back(']');
}

int resolveThreads(int threads, int64_t n) {
  if (threads <= 0) threads = (int)std::thread::hardware_concurrency();
  if (threads <= 0) threads = 1;
  return (int)std::max<int64_t>(1, std::min<int64_t>(threads, (n + 255) / 256));
}

void parseOne(std::string_view msg, bool tagged, FeatureBatch& b) {
  Cur c{msg.data(), msg.data() + msg.size()};
  skipWs(c);
  if (!c.more() || *c.p != '[') bad("message is not a JSON array");
  c.p++;
  char matrix = 0;
  std::string tmp;
  if (tagged) {
    readScalarAsString(c, tmp);
    if (tmp == "X") matrix = 'X';
    else if (tmp == "Y") matrix = 'Y';
    else matrix = '?';  // reported after the vector parses, where the reference's inner switch throws
    skipWs(c);
    if (!c.more() || *c.p != ',') bad("update has fewer than three elements");
    c.p++;
  }
  b.ids.emplace_back();
  readScalarAsString(c, b.ids.back());
  skipWs(c);
  if (!c.more() || *c.p != ',') bad("message has no vector element");
  c.p++;
  readFloatArray(c, b.values);
  if (matrix == '?') bad("matrix tag is neither \"X\" nor \"Y\"");
  bool hasOthers = false;
  skipWs(c);
  if (c.more() && *c.p == ',' && tagged) {
    c.p++;
    readStringArray(c, b.others, /*scalarsToString=*/true);
    hasOthers = true;
    skipWs(c);
  }
  while (c.more() && *c.p == ',') {  // further elements are read and ignored
    c.p++;
    skipValue(c);
    skipWs(c);
  }
  if (!c.more() || *c.p != ']') bad("expected ',' or ']'");
  b.matrix.push_back(matrix);
  b.hasOthers.push_back(hasOthers ? 1 : 0);
  b.vecBegin.push_back((int64_t)b.values.size());
  b.otherBegin.push_back((int64_t)b.others.size());
}

template <typename GetMessage>
std::vector<FeatureBatch> parseParts(GetMessage get, int64_t n, bool tagged, int threads) {
  const int T = resolveThreads(threads, n);
  std::vector<FeatureBatch> parts((size_t)T);
  auto work = [&](int t) {
    FeatureBatch& b = parts[(size_t)t];
    const int64_t lo = n * t / T, hi = n * (t + 1) / T;
    b.vecBegin.push_back(0);
    b.otherBegin.push_back(0);
    b.ids.reserve((size_t)(hi - lo));
    for (int64_t i = lo; i < hi; i++) {
      const size_t nIds = b.matrix.size();
      try {
        parseOne(get(i), tagged, b);
      } catch (const BadMessage& ex) {
        b.ids.resize(nIds);
        b.values.resize((size_t)b.vecBegin.back());
        b.others.resize((size_t)b.otherBegin.back());
        b.firstBad = i;
        b.firstBadWhy = ex.what();
        return;
      }
    }
  };
  if (T == 1) {
    work(0);
  } else {
    std::vector<std::thread> pool;
    for (int t = 0; t < T; t++) pool.emplace_back(work, t);
    for (auto& th : pool) th.join();
  }
  // everything after the first bad message is dropped
  for (size_t t = 0; t < parts.size(); t++) {
    if (parts[t].firstBad >= 0) { parts.resize(t + 1); break; }
  }
  return parts;
}

FeatureBatch mergeParts(std::vector<FeatureBatch>&& parts) {
  if (parts.size() == 1) return std::move(parts[0]);
  FeatureBatch out;
  out.vecBegin.push_back(0);
  out.otherBegin.push_back(0);
  for (auto& p : parts) {
    const int64_t v0 = (int64_t)out.values.size(), o0 = (int64_t)out.others.size();
    out.matrix.insert(out.matrix.end(), p.matrix.begin(), p.matrix.end());
    std::move(p.ids.begin(), p.ids.end(), std::back_inserter(out.ids));
    out.values.insert(out.values.end(), p.values.begin(), p.values.end());
    std::move(p.others.begin(), p.others.end(), std::back_inserter(out.others));
    out.hasOthers.insert(out.hasOthers.end(), p.hasOthers.begin(), p.hasOthers.end());
    for (size_t i = 1; i < p.vecBegin.size(); i++) out.vecBegin.push_back(v0 + p.vecBegin[i]);
    for (size_t i = 1; i < p.otherBegin.size(); i++) out.otherBegin.push_back(o0 + p.otherBegin[i]);
    if (p.firstBad >= 0) { out.firstBad = p.firstBad; out.firstBadWhy = p.firstBadWhy; }
  }
  return out;
}

// ---------------------------------------------------------------- commons-csv (RFC4180 + escape '\\')
constexpr int kNotMeta = -2;
int csvReadEscape(std::string_view s, size_t& i, char delimiter) {
  if (i >= s.size()) bad("EOF whilst processing escape sequence");
  char ch = s[i++];
  switch (ch) {
    case 'r': return '\r';
    case 'n': return '\n';
    case 't': return '\t';
    case 'b': return '\b';
    case 'f': return '\f';
    case '\r': case '\n': case '\f': case '\t': case '\b': return ch;
    default:
      if (ch == delimiter || ch == '\\' || ch == '"') return ch;
      return kNotMeta;
  }
}

// ---------------------------------------------------------------- tiny XML scanner for PMML
std::string xmlUnescape(std::string_view s) {
  std::string out;
  out.reserve(s.size());
  for (size_t i = 0; i < s.size();) {
    if (s[i] != '&') { out.push_back(s[i++]); continue; }
    size_t semi = s.find(';', i);
    if (semi == std::string_view::npos) bad("unterminated entity in PMML");
    std::string_view ent = s.substr(i + 1, semi - i - 1);
    if (ent == "lt") out.push_back('<');
    else if (ent == "gt") out.push_back('>');
    else if (ent == "amp") out.push_back('&');
    else if (ent == "quot") out.push_back('"');
    else if (ent == "apos") out.push_back('\'');
    else if (!ent.empty() && ent[0] == '#') {
      uint32_t cp = 0;
      if (ent.size() > 1 && (ent[1] == 'x' || ent[1] == 'X')) {
        for (size_t k = 2; k < ent.size(); k++) {
          int h = hexVal(ent[k]);
          if (h < 0) bad("bad character reference in PMML");
          cp = cp * 16 + (uint32_t)h;
        }
      } else {
        for (size_t k = 1; k < ent.size(); k++) {
          if (ent[k] < '0' || ent[k] > '9') bad("bad character reference in PMML");
          cp = cp * 10 + (uint32_t)(ent[k] - '0');
        }
      }
      appendUtf8(out, cp);
    } else {
      bad("unknown entity in PMML");
    }
    i = semi + 1;
  }
  return out;
}

std::string_view localName(std::string_view qname) {
  size_t colon = qname.find(':');
  return colon == std::string_view::npos ? qname : qname.substr(colon + 1);
}

double parseJavaDouble(const std::string& raw) {  // Double.valueOf(String)
  size_t b = 0, e = raw.size();
  while (b < e && (unsigned char)raw[b] <= ' ') b++;
  while (e > b && (unsigned char)raw[e - 1] <= ' ') e--;
  std::string s = raw.substr(b, e - b);
  if (s.empty()) bad("empty String where a number is required");  // NumberFormatException
  const char* p = s.c_str();
  bool neg = false;
  if (*p == '+' || *p == '-') { neg = *p == '-'; p++; }
  double v;
  if (std::strcmp(p, "NaN") == 0) v = std::numeric_limits<double>::quiet_NaN();
  else if (std::strcmp(p, "Infinity") == 0) v = std::numeric_limits<double>::infinity();
  else {
    const char* end = s.c_str() + s.size();
    if (end > p && (end[-1] == 'd' || end[-1] == 'D' || end[-1] == 'f' || end[-1] == 'F')) end--;
    if (end == p || !((*p >= '0' && *p <= '9') || *p == '.')) bad("For input string: \"" + raw + "\"");
    auto r = std::from_chars(p, end, v);
    if (r.ptr != end) bad("For input string: \"" + raw + "\"");
    if (r.ec != std::errc()) v = std::strtod(std::string(p, end).c_str(), nullptr);
  }
  return neg ? -v : v;
}

int64_t parseJavaLong(const std::string& s) {  // Long.valueOf(String): optional sign, digits only
  if (s.empty()) bad("For input string: \"\"");
  size_t i = (s[0] == '-' || s[0] == '+') ? 1 : 0;
  if (i == s.size()) bad("For input string: \"" + s + "\"");
  int64_t v = 0;
  auto r = std::from_chars(s.c_str() + (s[0] == '+' ? 1 : 0), s.c_str() + s.size(), v);
  if (r.ec != std::errc() || r.ptr != s.c_str() + s.size()) bad("For input string: \"" + s + "\"");
  return v;
}

}  // namespace

// ================================================================ TextUtils
namespace TextUtils {

std::vector<std::string> parseDelimited(std::string_view s, char delimiter) {
  std::vector<std::string> out;
  if (s.empty()) return {""};  // TextUtils.java:40,82-84 EMPTY_STRING
  size_t i = 0;
  for (;;) {
    std::string tok;
    bool endOfRecord = false;
    if (i < s.size() && s[i] == '"') {
      i++;
      for (;;) {
        if (i >= s.size()) bad("EOF reached before encapsulated token finished");
        char ch = s[i++];
        if (ch == '\\') {
          const char escaped = i < s.size() ? s[i] : 0;
          int u = csvReadEscape(s, i, delimiter);
          if (u == kNotMeta) { tok.push_back('\\'); tok.push_back(escaped); }
          else tok.push_back((char)u);
        } else if (ch == '"') {
          if (i < s.size() && s[i] == '"') { tok.push_back('"'); i++; continue; }
          for (;;) {  // up to the delimiter only whitespace may follow
            if (i >= s.size()) { endOfRecord = true; break; }
            char c2 = s[i++];
            if (c2 == delimiter) break;
            if (c2 == '\n' || c2 == '\r') { endOfRecord = true; break; }
            if (!(c2 == ' ' || c2 == '\t' || c2 == '\f' || c2 == '\v'))
              bad("invalid char between encapsulated token and delimiter");
          }
          break;
        } else {
          tok.push_back(ch);
        }
      }
    } else {
      for (;;) {
        if (i >= s.size()) { endOfRecord = true; break; }
        char ch = s[i++];
        if (ch == delimiter) break;
        if (ch == '\n' || ch == '\r') { endOfRecord = true; break; }
        if (ch == '\\') {
          const char escaped = i < s.size() ? s[i] : 0;
          int u = csvReadEscape(s, i, delimiter);
          if (u == kNotMeta) { tok.push_back('\\'); tok.push_back(escaped); }
          else tok.push_back((char)u);
        } else {
          tok.push_back(ch);
        }
      }
    }
    out.push_back(std::move(tok));
    if (endOfRecord) break;
    if (i >= s.size()) { out.emplace_back(); break; }  // trailing delimiter -> one more, empty, value
  }
  return out;
}

std::vector<std::string> parsePMMLDelimited(std::string_view delimited) {
  std::vector<std::string> out;
  for (auto& raw : parseDelimited(delimited, ' '))
    if (!raw.empty()) out.push_back(std::move(raw));
  return out;
}

std::vector<std::string> parseJSONArray(std::string_view json) {
  Cur c{json.data(), json.data() + json.size()};
  std::vector<std::string> out;
  skipWs(c);
  if (!c.more() || *c.p != '[') bad("Cannot deserialize String[] out of a non-array token");
  // String[] deserialisation coerces scalar tokens to their text; null stays null (reported as "")
  c.p++;
  skipWs(c);
  if (c.more() && *c.p == ']') return out;
  for (;;) {
    skipWs(c);
    out.emplace_back();
    if (c.more() && matchLiteral(c, "null")) {
      // a Java null element
    } else if (c.more() && (*c.p == '-' || (*c.p >= '0' && *c.p <= '9'))) {
      NumTok t = readNumberToken(c);
      out.back().assign(t.b, t.e);
    } else {
      readScalarAsString(c, out.back());
    }
    skipWs(c);
    if (!c.more()) bad("unexpected end of input");
    if (*c.p == ',') { c.p++; continue; }
    if (*c.p == ']') break;
    bad("expected ',' or ']'");
  }
  return out;
}

std::string joinJSON(const std::vector<std::string>& elements) {
  std::string out = "[";
  for (size_t i = 0; i < elements.size(); i++) {
    if (i) out.push_back(',');
    appendJSONString(out, elements[i]);
  }
  out.push_back(']');
  return out;
}

std::string floatToString(float v) { return javaToString(v); }
std::string doubleToString(double v) { return javaToString(v); }

}  // namespace TextUtils

// ================================================================ MLFunctions
namespace MLFunctions {

std::vector<std::string> parse(std::string_view line) {
  if (!line.empty() && line.front() == '[' && line.back() == ']') return TextUtils::parseJSONArray(line);
  return TextUtils::parseDelimited(line, ',');
}

int64_t toTimestamp(std::string_view line) {
  auto tokens = parse(line);
  if (tokens.size() < 4) throw std::out_of_range("ArrayIndexOutOfBoundsException: 3");
  return parseJavaLong(tokens[3]);
}

double sumWithNaN(const std::vector<double>& orderedStrengths) {
  double finalStrength = std::numeric_limits<double>::quiet_NaN();
  for (double strength : orderedStrengths) {
    if (std::isnan(finalStrength)) finalStrength = strength;
    else finalStrength += strength;
  }
  return finalStrength;
}

}  // namespace MLFunctions

// ================================================================ PMMLExtensions
PMMLExtensions PMMLExtensions::fromString(std::string_view xml) {
  PMMLExtensions out;
  int depth = 0;
  size_t i = 0;
  bool inExtension = false;     // inside a top-level <Extension> with content
  int extensionDepth = 0;
  std::string extName, extText;
  int extChildren = 0;
  bool extRecord = false;
  std::unordered_map<std::string, bool> seen;
  bool sawRoot = false;
  while (i < xml.size()) {
    size_t lt = xml.find('<', i);
    if (lt == std::string_view::npos) lt = xml.size();
    if (inExtension && depth == extensionDepth && lt > i) extText += xmlUnescape(xml.substr(i, lt - i));
    if (lt >= xml.size()) break;
    i = lt;
    if (xml.compare(i, 4, "<!--") == 0) {
      size_t end = xml.find("-->", i + 4);
      if (end == std::string_view::npos) bad("unterminated comment in PMML");
      i = end + 3;
    } else if (xml.compare(i, 9, "<![CDATA[") == 0) {
      size_t end = xml.find("]]>", i + 9);
      if (end == std::string_view::npos) bad("unterminated CDATA in PMML");
      if (inExtension && depth == extensionDepth) extText.append(xml.substr(i + 9, end - i - 9));
      i = end + 3;
    } else if (xml.compare(i, 2, "<?") == 0) {
      size_t end = xml.find("?>", i + 2);
      if (end == std::string_view::npos) bad("unterminated processing instruction in PMML");
      i = end + 2;
    } else if (xml.compare(i, 2, "<!") == 0) {
      size_t end = xml.find('>', i + 2);
      if (end == std::string_view::npos) bad("unterminated declaration in PMML");
      i = end + 1;
    } else if (xml.compare(i, 2, "</") == 0) {
      size_t end = xml.find('>', i + 2);
      if (end == std::string_view::npos) bad("unterminated end tag in PMML");
      if (inExtension && depth == extensionDepth) {
        if (extRecord) {
          if (extChildren == 0 && !extText.empty()) out.contents[extName] = TextUtils::parsePMMLDelimited(extText);
          else out.contents[extName] = {};
        }
        inExtension = false;
      }
      depth--;
      i = end + 1;
    } else {
      // start tag: name, attributes, optional "/"
      size_t p = i + 1;
      size_t nameEnd = p;
      while (nameEnd < xml.size() && !std::strchr(" \t\r\n/>", xml[nameEnd])) nameEnd++;
      std::string_view name = localName(xml.substr(p, nameEnd - p));
      std::unordered_map<std::string, std::string> attrs;
      p = nameEnd;
      bool selfClosed = false;
      for (;;) {
        while (p < xml.size() && std::strchr(" \t\r\n", xml[p])) p++;
        if (p >= xml.size()) bad("unterminated start tag in PMML");
        if (xml[p] == '>') { p++; break; }
        if (xml[p] == '/') {
          if (p + 1 >= xml.size() || xml[p + 1] != '>') bad("malformed empty-element tag in PMML");
          selfClosed = true;
          p += 2;
          break;
        }
        size_t eq = xml.find('=', p);
        if (eq == std::string_view::npos) bad("attribute without a value in PMML");
        std::string attrName(xml.substr(p, eq - p));
        while (!attrName.empty() && std::strchr(" \t\r\n", attrName.back())) attrName.pop_back();
        p = eq + 1;
        while (p < xml.size() && std::strchr(" \t\r\n", xml[p])) p++;
        if (p >= xml.size() || (xml[p] != '"' && xml[p] != '\'')) bad("unquoted attribute value in PMML");
        const char quote = xml[p++];
        size_t close = xml.find(quote, p);
        if (close == std::string_view::npos) bad("unterminated attribute value in PMML");
        attrs[std::string(localName(attrName))] = xmlUnescape(xml.substr(p, close - p));
        p = close + 1;
      }
      if (!sawRoot) {
        if (name != "PMML") bad("document element is not PMML");
        sawRoot = true;
      }
      if (inExtension && depth == extensionDepth) extChildren++;
      if (depth == 1 && name == "Extension" && !inExtension) {
        auto nameIt = attrs.find("name");
        const std::string en = nameIt == attrs.end() ? std::string() : nameIt->second;
        auto valueIt = attrs.find("value");
        // findFirst(): the first extension of a given name wins (AppPMMLUtils.java:63-64, 74-75)
        const bool first = seen.emplace(en, true).second;
        if (first && valueIt != attrs.end()) out.values[en] = valueIt->second;
        if (first && selfClosed) out.contents[en] = {};
        if (!selfClosed) {
          inExtension = true;
          extensionDepth = depth + 1;
          extName = en;
          extText.clear();
          extChildren = 0;
          extRecord = first;
        }
      }
      if (!selfClosed) depth++;
      i = p;
    }
  }
  if (!sawRoot) bad("no PMML document element");
  return out;
}

PMMLExtensions PMMLExtensions::fromUpdateKeyMessage(const std::string& key, const std::string& message) {
  if (key == "MODEL") return fromString(message);
  if (key == "MODEL-REF") {
    std::string path = message;
    if (path.rfind("file:", 0) == 0) {
      path = path.substr(5);
      while (path.size() > 1 && path[0] == '/' && path[1] == '/') path.erase(0, 1);
    }
    std::ifstream in(path, std::ios::binary);
    if (!in) bad("cannot open model file " + message);
    std::stringstream ss;
    ss << in.rdbuf();
    return fromString(ss.str());
  }
  throw std::invalid_argument("Unknown key " + key);
}

// ================================================================ feature-vector messages
FeatureBatch parseFeatureMessages(const char* text, const int64_t* begin, int64_t n, bool tagged, int threads) {
  auto get = [&](int64_t i) { return std::string_view(text + begin[i], (size_t)(begin[i + 1] - begin[i])); };
  return mergeParts(parseParts(get, n, tagged, threads));
}

std::vector<FeatureBatch> parseFeatureMessageParts(const char* text, const int64_t* begin, int64_t n, bool tagged,
                                                   int threads) {
  auto get = [&](int64_t i) { return std::string_view(text + begin[i], (size_t)(begin[i + 1] - begin[i])); };
  return parseParts(get, n, tagged, threads);
}

void parseFeatureMessage(std::string_view message, bool tagged, char* matrix, std::string* id,
                         std::vector<float>* vector, bool* hasOthers, std::vector<std::string>* others) {
  FeatureBatch b;
  b.vecBegin.push_back(0);
  b.otherBegin.push_back(0);
  parseOne(message, tagged, b);
  if (matrix) *matrix = b.matrix[0];
  if (id) *id = std::move(b.ids[0]);
  if (vector) *vector = std::move(b.values);
  if (hasOthers) *hasOthers = b.hasOthers[0] != 0;
  if (others) *others = std::move(b.others);
}

std::string toUpdateJSON(char matrix, const std::string& id, const float* vector, int features,
                         const std::string* other) {
  std::string out = "[";
  appendJSONString(out, std::string(1, matrix));
  out.push_back(',');
  appendJSONString(out, id);
  out.push_back(',');
  appendJSONFloats(out, vector, features);
  if (other) {
    out += ",[";
    appendJSONString(out, *other);
    out.push_back(']');
  }
  out.push_back(']');
  return out;
}

std::string toFeatureLine(const std::string& id, const float* vector, int features) {
  std::string out = "[";
  appendJSONString(out, id);
  out.push_back(',');
  appendJSONFloats(out, vector, features);
  out.push_back(']');
  return out;
}

// ================================================================ ALSServingModelManager
ALSServingModelManager::ALSServingModelManager(double sampleRate, int numCores, int device, int parseThreads,
                                               int64_t maxBatchRows)
    : sampleRate_(sampleRate), numCores_(numCores), device_(device), parseThreads_(parseThreads),
      maxBatchRows_(std::max<int64_t>(1, maxBatchRows)) {
  if (!(sampleRate > 0.0 && sampleRate <= 1.0)) throw std::invalid_argument("sampleRate must be in (0,1]");  // :60
}

void ALSServingModelManager::applyBatch(const FeatureBatch& b, char forceMatrix) {
  ALSServingModel& model = *model_;
  const int F = model.getFeatures();
  const size_t n = b.size();
  // Item vectors of one batch go to the device in runs; a length mismatch is the reference's
  // checkArgument(vector.length == features) and stops the stream at that message.
  size_t i = 0;
  std::vector<float> gathered;
  while (i < n) {
    const char mx = forceMatrix ? forceMatrix : b.matrix[i];
    if (b.vecLen(i) != F) {
      throw std::invalid_argument("vector.length != features for id " + b.ids[i]);
    }
    if (mx == 'Y') {
      size_t j = i;
      while (j < n && (forceMatrix ? forceMatrix : b.matrix[j]) == 'Y' && b.vecLen(j) == F) j++;
      model.setItemVectors(b.ids.data() + i, b.vec(i), j - i);  // vectors of a run are contiguous
      i = j;
    } else {
      size_t j = i;
      while (j < n && (forceMatrix ? forceMatrix : b.matrix[j]) == 'X' && b.vecLen(j) == F) j++;
      model.setUserVectors(b.ids.data() + i, b.vec(i), j - i);
      // known items of the run in one call (only messages that carry a fourth element, :83-87)
      std::vector<std::string> users;
      std::vector<int64_t> begin{0};
      std::vector<std::string> items;
      for (size_t k = i; k < j; k++) {
        if (!b.hasOthers[k]) continue;
        users.push_back(b.ids[k]);
        items.insert(items.end(), b.others.begin() + b.otherBegin[k], b.others.begin() + b.otherBegin[k + 1]);
        begin.push_back((int64_t)items.size());
      }
      if (!users.empty()) model.addKnownItems(users.data(), begin.data(), items.data(), users.size());
      i = j;
    }
  }
}

namespace {
// Parses sub-batch i+1 on the worker threads while the caller applies sub-batch i (bookkeeping and the
// device upsert are serial: one updater).
template <typename Apply>
void pipelinedParse(const char* text, const int64_t* begin, int64_t n, bool tagged, int threads, int64_t maxBatchRows,
                    Apply apply) {
  auto parseRange = [=](int64_t lo) {
    const int64_t hi = std::min(n, lo + maxBatchRows);
    return parseFeatureMessageParts(text, begin + lo, hi - lo, tagged, threads);
  };
  if (n <= 0) return;
  std::future<std::vector<FeatureBatch>> next = std::async(std::launch::async, parseRange, (int64_t)0);
  for (int64_t lo = 0; lo < n; lo += maxBatchRows) {
    std::vector<FeatureBatch> parts = next.get();
    if (lo + maxBatchRows < n) next = std::async(std::launch::async, parseRange, lo + maxBatchRows);
    try {
      apply(parts, lo);
    } catch (...) {
      if (next.valid()) next.wait();  // do not leave a parse running over the caller's buffers
      throw;
    }
  }
}
}  // namespace

void ALSServingModelManager::consumeUpdates(const char* text, const int64_t* begin, int64_t n) {
  if (!model_) return;  // :73-75 no model to interpret with yet
  pipelinedParse(text, begin, n, /*tagged=*/true, parseThreads_, maxBatchRows_,
                 [&](std::vector<FeatureBatch>& parts, int64_t lo) {
                   for (auto& p : parts) {
                     applyBatch(p, 0);
                     if (p.firstBad >= 0)
                       throw BadMessage("Bad message #" + std::to_string(lo + p.firstBad) + ": " + p.firstBadWhy);
                   }
                 });
}

void ALSServingModelManager::loadFeatureLines(char matrix, const char* text, const int64_t* begin, int64_t n) {
  if (!model_) throw std::logic_error("no model: a MODEL message must come first");
  if (matrix != 'X' && matrix != 'Y') throw std::invalid_argument("matrix must be 'X' or 'Y'");
  pipelinedParse(text, begin, n, /*tagged=*/false, parseThreads_, maxBatchRows_,
                 [&](std::vector<FeatureBatch>& parts, int64_t lo) {
                   for (auto& p : parts) {
                     applyBatch(p, matrix);
                     if (p.firstBad >= 0)
                       throw BadMessage("Bad line #" + std::to_string(lo + p.firstBad) + ": " + p.firstBadWhy);
                   }
                 });
}

namespace {
struct ModelHeader {
  int features;
  bool implicit;
  std::vector<std::string> xids, yids;
};
ModelHeader readModelHeader(const std::string& key, const std::string& message) {
  PMMLExtensions pmml = PMMLExtensions::fromUpdateKeyMessage(key, message);
  ModelHeader h;
  auto f = pmml.values.find("features");
  if (f == pmml.values.end()) bad("NumberFormatException: null (no \"features\" extension)");
  h.features = (int)parseJavaLong(f->second);
  auto imp = pmml.values.find("implicit");
  std::string iv = imp == pmml.values.end() ? std::string() : imp->second;
  std::transform(iv.begin(), iv.end(), iv.begin(), [](unsigned char ch) { return (char)std::tolower(ch); });
  h.implicit = iv == "true";  // Boolean.valueOf
  auto x = pmml.contents.find("XIDs");
  auto y = pmml.contents.find("YIDs");
  // new HashSet<>(null) is a NullPointerException in the reference
  if (x == pmml.contents.end() || y == pmml.contents.end()) bad("model has no XIDs / YIDs extension content");
  h.xids = x->second;
  h.yids = y->second;
  return h;
}

// Runs of consecutive "UP" messages are handed to `ups` as [first, last); anything else to `other`.
void forEachRun(const std::vector<KeyMessage>& updates, const std::function<void(size_t, size_t)>& ups,
                const std::function<void(const KeyMessage&)>& other) {
  size_t i = 0;
  while (i < updates.size()) {
    if (updates[i].key == "UP") {
      size_t j = i;
      while (j < updates.size() && updates[j].key == "UP") j++;
      ups(i, j);
      i = j;
    } else {
      other(updates[i]);
      i++;
    }
  }
}
}  // namespace

void ALSServingModelManager::consume(const std::vector<KeyMessage>& updates) {
  forEachRun(
      updates,
      [&](size_t first, size_t last) {
        if (!model_) return;
        for (size_t lo = first; lo < last; lo += (size_t)maxBatchRows_) {
          const size_t hi = std::min(last, lo + (size_t)maxBatchRows_);
          auto get = [&](int64_t i) { return std::string_view(updates[lo + (size_t)i].message); };
          auto parts = parseParts(get, (int64_t)(hi - lo), true, parseThreads_);
          for (auto& p : parts) {
            applyBatch(p, 0);
            if (p.firstBad >= 0)
              throw BadMessage("Bad message: " + updates[lo + (size_t)p.firstBad].message + " (" + p.firstBadWhy + ")");
          }
        }
      },
      [&](const KeyMessage& km) {
        if (km.key != "MODEL" && km.key != "MODEL-REF")
          throw std::invalid_argument("Bad message: " + km.key + "," + km.message);  // :125-126
        ModelHeader h = readModelHeader(km.key, km.message);
        if (!model_ || h.features != model_->getFeatures())  // :110-113
          model_ = std::make_unique<ALSServingModel>(h.features, h.implicit, sampleRate_, numCores_, device_);
        model_->retainRecentAndKnownItems(h.xids, h.yids);  // :119-121
        model_->retainRecentAndUserIDs(h.xids);
        model_->retainRecentAndItemIDs(h.yids);
      });
}

// ================================================================ ALSSpeedModelManager
ALSSpeedModelManager::ALSSpeedModelManager(bool noKnownItems, double minModelLoadFraction, int device, int parseThreads)
    : noKnownItems_(noKnownItems), minModelLoadFraction_(minModelLoadFraction), device_(device),
      parseThreads_(parseThreads) {
  if (!(minModelLoadFraction >= 0.0 && minModelLoadFraction <= 1.0))
    throw std::invalid_argument("minModelLoadFraction must be in [0,1]");  // :62
}

void ALSSpeedModelManager::consume(const std::vector<KeyMessage>& updates) {
  forEachRun(
      updates,
      [&](size_t first, size_t last) {
        if (!model_) return;  // :75-77
        auto get = [&](int64_t i) { return std::string_view(updates[first + (size_t)i].message); };
        auto parts = parseParts(get, (int64_t)(last - first), true, parseThreads_);
        const int F = model_->getFeatures();
        for (auto& p : parts) {
          size_t i = 0;
          const size_t n = p.size();
          while (i < n) {
            if (p.vecLen(i) != F) throw std::invalid_argument("vector.length != features for id " + p.ids[i]);
            size_t j = i;
            while (j < n && p.matrix[j] == p.matrix[i] && p.vecLen(j) == F) j++;
            model_->setVectors(p.matrix[i], p.ids.data() + i, p.vec(i), j - i);
            i = j;
          }
          if (p.firstBad >= 0)
            throw BadMessage("Bad message: " + updates[first + (size_t)p.firstBad].message + " (" + p.firstBadWhy + ")");
        }
      },
      [&](const KeyMessage& km) {
        if (km.key != "MODEL" && km.key != "MODEL-REF")
          throw std::invalid_argument("Bad message: " + km.key + "," + km.message);  // :128-129
        ModelHeader h = readModelHeader(km.key, km.message);
        if (!model_ || h.features != model_->getFeatures())  // :113-116
          model_ = std::make_unique<ALSSpeedModel>(h.features, h.implicit, device_);
        model_->retainRecentAndUserIDs(h.xids);  // :122-123
        model_->retainRecentAndItemIDs(h.yids);
      });
}

std::vector<UserItemStrength> ALSSpeedModelManager::aggregate(const std::vector<std::string>& lines,
                                                              bool implicit) const {
  struct Row {
    int64_t timestamp;
    size_t index;
    std::string user, item;
    double strength;
  };
  std::vector<Row> rows(lines.size());
  const int T = resolveThreads(parseThreads_, (int64_t)lines.size());
  std::atomic<int64_t> firstBad{-1};
  std::vector<std::string> why((size_t)T);
  auto work = [&](int t) {
    const size_t lo = lines.size() * (size_t)t / (size_t)T, hi = lines.size() * (size_t)(t + 1) / (size_t)T;
    for (size_t i = lo; i < hi; i++) {
      try {
        auto tokens = MLFunctions::parse(lines[i]);
        if (tokens.size() < 4) throw std::out_of_range("ArrayIndexOutOfBoundsException");
        rows[i] = Row{parseJavaLong(tokens[3]), i, std::move(tokens[0]), std::move(tokens[1]), parseJavaDouble(tokens[2])};
      } catch (const std::exception& ex) {
        int64_t expected = -1;
        if (firstBad.compare_exchange_strong(expected, (int64_t)i)) why[(size_t)t] = ex.what();
        return;
      }
    }
  };
  if (T == 1) work(0);
  else {
    std::vector<std::thread> pool;
    for (int t = 0; t < T; t++) pool.emplace_back(work, t);
    for (auto& th : pool) th.join();
  }
  if (firstBad.load() >= 0) {
    std::string w;
    for (auto& s : why) if (!s.empty()) w = s;
    throw BadMessage("Bad input: " + lines[(size_t)firstBad.load()] + " (" + w + ")");  // :153-156
  }
  std::stable_sort(rows.begin(), rows.end(), [](const Row& a, const Row& b) { return a.timestamp < b.timestamp; });  // :142-143
  if (model_ && model_->handle() != nullptr && !rows.empty()) {
    // With a model (hence a GPU) at hand the grouping and the ordered reduction run on the device
    // (oryx_aggregate_interactions): ids are interned to batch-local integers in order of appearance, so a
    // user or item the model has not seen yet takes part like any other.
    std::unordered_map<std::string, int32_t> uid, iid;
    std::vector<const std::string*> unames, inames;
    std::vector<int32_t> u(rows.size()), it(rows.size());
    std::vector<double> st(rows.size());
    std::vector<int64_t> ts(rows.size());
    for (size_t k = 0; k < rows.size(); k++) {
      auto a = uid.emplace(rows[k].user, (int32_t)uid.size());
      if (a.second) unames.push_back(&a.first->first);
      auto b = iid.emplace(rows[k].item, (int32_t)iid.size());
      if (b.second) inames.push_back(&b.first->first);
      u[k] = a.first->second;
      it[k] = b.first->second;
      st[k] = rows[k].strength;
      ts[k] = rows[k].timestamp;
    }
    std::vector<int32_t> ou(rows.size()), oi(rows.size());
    std::vector<double> os(rows.size());
    int64_t kept = 0;
    const int32_t status = oryx_aggregate_interactions(model_->handle(), u.data(), it.data(), st.data(), ts.data(),
                                                       (int64_t)rows.size(), implicit ? 1 : 0, ou.data(), oi.data(), os.data(), &kept);
    if (status != ORYX_OK) {
      char msg[512] = {0};
      oryx_last_error(msg, sizeof(msg));
      throw std::runtime_error(std::string("oryx_aggregate_interactions: ") + msg);
    }
    std::vector<UserItemStrength> result;
    result.reserve((size_t)kept);
    for (int64_t k = 0; k < kept; k++) result.push_back(UserItemStrength{*unames[(size_t)ou[(size_t)k]], *inames[(size_t)oi[(size_t)k]], os[(size_t)k]});
    return result;
  }
  // no model yet (nothing to fold into either): the same reduction on the host
  // (user,item) -> slot in first-appearance order; the key joins the two ids with a byte that cannot
  // appear in either after parsing a line
  std::unordered_map<std::string, size_t> slot;
  std::vector<UserItemStrength> out;
  slot.reserve(rows.size());
  for (auto& r : rows) {
    std::string key;
    key.reserve(r.user.size() + r.item.size() + 9);
    const uint64_t ulen = r.user.size();
    key.append(reinterpret_cast<const char*>(&ulen), sizeof(ulen));
    key += r.user;
    key += r.item;
    auto it = slot.find(key);
    if (it == slot.end()) {
      slot.emplace(std::move(key), out.size());
      out.push_back(UserItemStrength{std::move(r.user), std::move(r.item), r.strength});
    } else if (implicit) {
      double& acc = out[it->second].strength;  // SUM_WITH_NAN, MLFunctions.java:67-78
      if (std::isnan(acc)) acc = r.strength;
      else acc += r.strength;
    } else {
      out[it->second].strength = r.strength;  // foldByKey(NaN, (current, next) -> next) :164
    }
  }
  std::vector<UserItemStrength> kept;
  kept.reserve(out.size());
  for (auto& u : out) {
    if (std::isnan(u.strength)) continue;           // :167-168
    u.strength = (double)(float)u.strength;         // strength.floatValue() :171, widened again at :184
    kept.push_back(std::move(u));
  }
  return kept;
}

std::vector<std::string> ALSSpeedModelManager::buildUpdates(const std::vector<std::string>& lines) {
  std::vector<std::string> result;
  if (!model_ || model_->getFractionLoaded() < minModelLoadFraction_) return result;  // :136-138
  auto input = aggregate(lines, model_->isImplicit());
  std::vector<Update> updates;
  try {
    updates = model_->buildUpdates(input);
  } catch (const SingularMatrixSolverException&) {
    return result;  // :177-179
  }
  result.resize(updates.size());
  const int F = model_->getFeatures();
  const int T = resolveThreads(parseThreads_, (int64_t)updates.size());
  auto work = [&](int t) {
    const size_t lo = updates.size() * (size_t)t / (size_t)T, hi = updates.size() * (size_t)(t + 1) / (size_t)T;
    for (size_t i = lo; i < hi; i++) {
      const Update& u = updates[i];
      result[i] = toUpdateJSON(u.matrix, u.id, u.vector.data(), F, noKnownItems_ ? nullptr : &u.other);
    }
  };
  if (T == 1) work(0);
  else {
    std::vector<std::thread> pool;
    for (int t = 0; t < T; t++) pool.emplace_back(work, t);
    for (auto& th : pool) th.join();
  }
  return result;
}

}  // namespace host
}  // namespace oryx
