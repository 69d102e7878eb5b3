// Byte-level logic of the response parser and the Countdown scorer, as pure host/device functions.
// The kernels in parse.cu call these; tests/test_hostsim.py runs the same code on the host against the oracle.
//
// Follows ContextManager._parse_response (ragen/llm_agent/ctx_manager.py:148-173), EnvStateManager.
// _extract_map_valid_actions (ragen/llm_agent/es_manager.py:230-240) and CountdownEnv.compute_reward /
// check_format / check_correctness (ragen/env/countdown/env.py:9-21, 69-78).  Text is UTF-8.
#pragma once
#include <math.h>
#include <stdint.h>

#ifndef RAGEN_HD
#ifdef __CUDACC__
#define RAGEN_HD __host__ __device__ __forceinline__
#else
#define RAGEN_HD inline
#endif
#endif

namespace parse {

// ---- Unicode whitespace (str.isspace / regex \s on str): U+0009-000D, 001C-001F, 0020, 0085, 00A0, 1680,
// 2000-200A, 2028, 2029, 202F, 205F, 3000 ------------------------------------------------------------------
// length in bytes of the whitespace character starting at p (0 if none)
RAGEN_HD int space_at(const uint8_t* p, const uint8_t* end) {
  const int c = p[0];
  if (c < 0x80) return ((c >= 0x09 && c <= 0x0d) || (c >= 0x1c && c <= 0x20)) ? 1 : 0;
  if (c == 0xc2) return (p + 1 < end && (p[1] == 0x85 || p[1] == 0xa0)) ? 2 : 0;
  if (p + 2 >= end) return 0;
  const int d = p[1], e = p[2];
  if (c == 0xe1) return (d == 0x9a && e == 0x80) ? 3 : 0;
  if (c == 0xe2) {
    if (d == 0x80) return ((e >= 0x80 && e <= 0x8a) || e == 0xa8 || e == 0xa9 || e == 0xaf) ? 3 : 0;
    if (d == 0x81) return (e == 0x9f) ? 3 : 0;
    return 0;
  }
  if (c == 0xe3) return (d == 0x80 && e == 0x80) ? 3 : 0;
  return 0;
}
// length in bytes of the whitespace character ENDING just before p (0 if none); begin <= p
RAGEN_HD int space_before(const uint8_t* begin, const uint8_t* p) {
  if (p <= begin) return 0;
  const int c = p[-1];
  if (c < 0x80) return ((c >= 0x09 && c <= 0x0d) || (c >= 0x1c && c <= 0x20)) ? 1 : 0;
  if (p - begin >= 2 && p[-2] == 0xc2) return (c == 0x85 || c == 0xa0) ? 2 : 0;
  if (p - begin >= 3) return space_at(p - 3, p) == 3 ? 3 : 0;
  return 0;
}
// str.strip(): shrink [s, e) from both ends
RAGEN_HD void strip(const uint8_t* base, int& s, int& e) {
  for (;;) {
    if (s >= e) break;
    const int k = space_at(base + s, base + e);
    if (!k) break;
    s += k;
  }
  for (;;) {
    if (s >= e) break;
    const int k = space_before(base + s, base + e);
    if (!k) break;
    e -= k;
  }
}

// ---- the six special strings (ctx_manager.py:94), all starting with '<' ------------------------------------
enum Tag { T_NONE = 0, T_THINK = 1, T_THINK_END = 2, T_ANSWER = 3, T_ANSWER_END = 4, T_IM_START = 5, T_IM_END = 6 };

RAGEN_HD int tag_len(int tag) {
  return tag == T_THINK ? 7 : tag == T_THINK_END ? 8 : tag == T_ANSWER ? 8 : tag == T_ANSWER_END ? 9 : tag == T_IM_START ? 12
                                                                                                 : tag == T_IM_END   ? 10
                                                                                                                     : 0;
}
RAGEN_HD bool bytes_eq(const uint8_t* p, const uint8_t* end, const char* lit, int n) {
  if (end - p < n) return false;
  for (int i = 0; i < n; ++i)
    if (p[i] != (uint8_t)lit[i]) return false;
  return true;
}
// which special string starts at p (p[0] is known to be '<')
RAGEN_HD int classify_tag(const uint8_t* p, const uint8_t* end) {
  if (end - p < 7) return T_NONE;
  const int c = p[1];
  if (c == 't') return bytes_eq(p, end, "<think>", 7) ? T_THINK : T_NONE;
  if (c == 'a') return bytes_eq(p, end, "<answer>", 8) ? T_ANSWER : T_NONE;
  if (c == '/') {
    if (p[2] == 't') return bytes_eq(p, end, "</think>", 8) ? T_THINK_END : T_NONE;
    return bytes_eq(p, end, "</answer>", 9) ? T_ANSWER_END : T_NONE;
  }
  if (c == '|') {
    if (bytes_eq(p, end, "<|im_start|>", 12)) return T_IM_START;
    return bytes_eq(p, end, "<|im_end|>", 10) ? T_IM_END : T_NONE;
  }
  return T_NONE;
}

// ---- re.search(r'<think>(.*?)</think>\s*<answer>(.*?)</answer>', text, re.DOTALL) as a state machine over the tag
// occurrences fed in text order (answer-only pattern when enable_think is false).
//
// Leftmost match with lazy groups: it starts at the FIRST <think>; group 1 ends at the first </think> that is
// followed by \s*<answer>; group 2 ends at the first </answer> after that.  If that fails no later start can
// succeed (any match from a later <think> is also a match from the first one).
struct Matcher {
  int state;  // 0 want <think> | 1 want </think> | 2 want </answer> | 3 matched | 4 want <answer> (answer-only)
  int think_s, think_e, ans_s, ans_e;
  int skip_until;  // occurrences starting before this offset lie inside an already consumed literal
};
RAGEN_HD void matcher_init(Matcher& m, bool enable_think) {
  m.state = enable_think ? 0 : 4;
  m.think_s = m.think_e = m.ans_s = m.ans_e = 0;
  m.skip_until = 0;
}
// feed one tag occurrence (tag at byte offset pos of text[0, n)); returns true once matched
RAGEN_HD bool matcher_feed(Matcher& m, const uint8_t* text, int n, int pos, int tag) {
  if (m.state == 3 || pos < m.skip_until) return m.state == 3;
  if (m.state == 0) {
    if (tag == T_THINK) {
      m.think_s = pos + 7;
      m.skip_until = pos + 7;
      m.state = 1;
    }
  } else if (m.state == 1) {
    if (tag == T_THINK_END) {
      int q = pos + 8;
      for (;;) {  // \s*
        if (q >= n) break;
        const int k = space_at(text + q, text + n);
        if (!k) break;
        q += k;
      }
      if (q < n && text[q] == '<' && classify_tag(text + q, text + n) == T_ANSWER) {
        m.think_e = pos;
        m.ans_s = q + 8;
        m.skip_until = q + 8;
        m.state = 2;
      }
    }
  } else if (m.state == 4) {
    if (tag == T_ANSWER) {
      m.ans_s = pos + 8;
      m.skip_until = pos + 8;
      m.state = 2;
    }
  } else if (m.state == 2) {
    if (tag == T_ANSWER_END) {
      m.ans_e = pos;
      m.state = 3;
    }
  }
  return m.state == 3;
}

// ---- "for special_token in list: content = content.replace(special_token, '').strip()" (ctx_manager.py:162-164) ----
// In place on buf[s, e); returns the new [s, e).  Six sequential passes, because removing one string can splice a
// later one together (e.g. "</ans<think>wer>").  str.replace scans left to right without overlaps.
RAGEN_HD void clean_content(uint8_t* buf, int& s, int& e) {
  for (int tag = T_THINK; tag <= T_IM_END; ++tag) {
    const int L = tag_len(tag);
    int r = s, w = s;
    while (r < e) {
      if (buf[r] == '<' && e - r >= L && classify_tag(buf + r, buf + e) == tag) {
        r += L;
      } else {
        buf[w++] = buf[r++];
      }
    }
    e = w;
    strip(buf, s, e);
  }
}
// does any special string start inside text[s, e)?  (then the slow path above is needed)
RAGEN_HD bool has_special(const uint8_t* text, int s, int e) {
  for (int i = s; i < e; ++i)
    if (text[i] == '<' && classify_tag(text + i, text + e) != T_NONE) return true;
  return false;
}

// ---- "[a.strip() for a in content.split(sep) if a.strip()]" then "[:max_actions]" (ctx_manager.py:166-171) --------
// spans[2k], spans[2k+1] = start, length of action k (k < max_actions); returns the number of non-empty actions
// found in total (so "> max_actions" tells the caller that the answer text is re-joined from the kept actions).
RAGEN_HD int split_actions(const uint8_t* buf, int s, int e, const uint8_t* sep, int sep_len, int max_actions, int* spans) {
  int count = 0;
  int piece = s;
  int i = s;
  for (;;) {
    bool at_sep = false;
    if (i >= e) {
      at_sep = true;
    } else if (sep_len > 0 && e - i >= sep_len) {
      at_sep = true;
      for (int k = 0; k < sep_len; ++k)
        if (buf[i + k] != sep[k]) {
          at_sep = false;
          break;
        }
    }
    if (at_sep) {
      int a = piece, b = i;
      strip(buf, a, b);
      if (b > a) {
        if (count < max_actions) {
          spans[2 * count] = a;
          spans[2 * count + 1] = b - a;
        }
        ++count;
      }
      if (i >= e) break;
      i += sep_len;
      piece = i;
    } else {
      ++i;
    }
  }
  return count;
}

// ---- rev_action_lookup[action.lower()] (es_manager.py:236-239) -----------------------------------------------------
// Table: n_entries names back to back in `names` (name k = names[name_off[k], name_off[k+1])), codes[k] the action id.
// Names are compared case-insensitively over ASCII letters (both sides are lower-cased by the reference); bytes >= 0x80
// must be equal.  With duplicate names the LAST entry wins (dict construction order).  Returns the code or -1.
RAGEN_HD int ascii_lower(int c) { return (c >= 'A' && c <= 'Z') ? c + 32 : c; }
RAGEN_HD int lookup_action(const uint8_t* act, int len, const uint8_t* names, const int32_t* name_off, const uint8_t* codes,
                           int first, int n_entries) {
  int found = -1;
  for (int k = first; k < first + n_entries; ++k) {
    const int a = name_off[k], b = name_off[k + 1];
    if (b - a != len) continue;
    bool same = true;
    for (int i = 0; i < len; ++i)
      if (ascii_lower(act[i]) != ascii_lower(names[a + i])) {
        same = false;
        break;
      }
    if (same) found = codes[k];
  }
  return found;
}

// ---- tail of the parse: split the cleaned answer, keep max_actions, map through the env's lookup table ------------
constexpr int MAX_ACTIONS = 16;
struct ActionTable {
  const uint8_t* names;
  const int32_t* name_off;
  const uint8_t* codes;
  int first, count;  // count < 0: env without action_lookup
};
struct ParseTail {
  int total, kept, nvalid;
  int spans[2 * MAX_ACTIONS];
  uint8_t codes[MAX_ACTIONS];
};
RAGEN_HD void parse_finish(const uint8_t* abuf, int ans_s, int ans_e, const uint8_t* sep, int sep_len, int max_actions,
                           const ActionTable& tab, ParseTail& out) {
  out.total = split_actions(abuf, ans_s, ans_e, sep, sep_len, max_actions, out.spans);
  out.kept = out.total < max_actions ? out.total : max_actions;
  out.nvalid = 0;
  if (tab.count >= 0)
    for (int k = 0; k < out.kept; ++k) {
      const int c = lookup_action(abuf + out.spans[2 * k], out.spans[2 * k + 1], tab.names, tab.name_off, tab.codes, tab.first,
                                  tab.count);
      if (c >= 0) out.codes[out.nvalid++] = (uint8_t)c;
    }
}

// ---- Countdown (ragen/env/countdown/env.py:9-21, 69-78) -------------------------------------------------------------
// check_format: the integers of all ASCII digit runs, as a multiset, must equal nums.
// check_correctness: eval(expr) == target within 1e-5.  The device evaluator covers this grammar of Python
// expressions: decimal integers (leading zeros only for zero, as Python 3), float literals (digits '.' digits, optional
// exponent; up to 15 significant digits), binary + - * / (true division) // (floor division) % ** (right associative,
// binding tighter than a unary sign on its left), unary + -, parentheses, blanks (space, tab, form feed; newlines inside
// parentheses or trailing), with Python's int / float result types.  Anything else Python might accept (names,
// comparisons, bit operators, complex results, digit separators, integers beyond 2^61 ...) is reported through
// `unsupported` and scored as an exception; callers surface that flag (CountdownEnv warns), they do not re-evaluate.
constexpr int CD_MAX_NUMS = 8;
constexpr int CD_STACK = 32;

struct CdValue {
  bool is_int;
  long long i;
  double d;
};
RAGEN_HD double cd_as_double(const CdValue& v) { return v.is_int ? (double)v.i : v.d; }
RAGEN_HD bool cd_finite(double x) { return x - x == 0.0; }
RAGEN_HD void cd_set_d(CdValue& out, double v) {
  out.is_int = false;
  out.i = 0;
  out.d = v;
}

// checked int64 multiply within |r| <= 2^61
RAGEN_HD bool cd_mul(long long a, long long b, long long& r) {
  bool ovf;
#if defined(__CUDA_ARCH__)
  r = (long long)((unsigned long long)a * (unsigned long long)b);
  ovf = (a != 0 && (r / a != b || (a == -1 && b == (-9223372036854775807LL - 1))));
#else
  ovf = __builtin_mul_overflow(a, b, &r);
#endif
  return !(ovf || r > (1LL << 61) || r < -(1LL << 61));
}

// CPython float_divmod (Objects/floatobject.c): floor quotient and the remainder with the divisor's sign; wx != 0
RAGEN_HD void cd_float_divmod(double vx, double wx, double& floordiv, double& mod) {
  mod = fmod(vx, wx);
  double div = (vx - mod) / wx;
  if (mod != 0.0) {
    if ((wx < 0) != (mod < 0)) {
      mod += wx;
      div -= 1.0;
    }
  } else {
    mod = copysign(0.0, wx);
  }
  if (div != 0.0) {
    floordiv = floor(div);
    if (div - floordiv > 0.5) floordiv += 1.0;
  } else {
    floordiv = copysign(0.0, vx / wx);
  }
}

// applies a binary operator ('f' = //, 'p' = **); false when Python raises (ZeroDivisionError, OverflowError) or the
// value leaves the covered range (then `unsupported` is set as well)
RAGEN_HD bool cd_apply(char op, const CdValue& a, const CdValue& b, CdValue& out, bool& unsupported) {
  const bool ints = a.is_int && b.is_int;
  if ((a.is_int && (a.i > (1LL << 53) || a.i < -(1LL << 53))) || (b.is_int && (b.i > (1LL << 53) || b.i < -(1LL << 53)))) {
    if (!ints || op == '/' || (op == 'p' && b.i < 0)) {  // int -> float of such magnitudes: Python rounds the exact quotient
      unsupported = true;
      return false;
    }
  }
  if (op == '/') {
    const double den = cd_as_double(b);
    if (den == 0.0) return false;
    cd_set_d(out, cd_as_double(a) / den);
    return true;
  }
  if (op == 'f' || op == '%') {
    if (ints) {
      if (b.i == 0) return false;
      long long q = a.i / b.i, r = a.i % b.i;
      if (r != 0 && ((r < 0) != (b.i < 0))) {
        q -= 1;
        r += b.i;
      }
      out.is_int = true;
      out.i = (op == 'f') ? q : r;
      out.d = 0.0;
      return true;
    }
    const double x = cd_as_double(a), y = cd_as_double(b);
    if (y == 0.0) return false;
    double fd, md;
    cd_float_divmod(x, y, fd, md);
    cd_set_d(out, (op == 'f') ? fd : md);
    return true;
  }
  if (op == 'p') {
    if (ints && b.i >= 0) {  // int ** non-negative int stays an int
      long long base = a.i, r = 1;
      long long e = b.i;
      if (base == 0 || base == 1) {
        r = (e == 0) ? 1 : base;
      } else if (base == -1) {
        r = (e & 1) ? -1 : 1;
      } else {
        if (e > 62) {
          unsupported = true;  // beyond int64: Python would carry on with big integers
          return false;
        }
        for (long long k = 0; k < e; ++k) {
          if (!cd_mul(r, base, r)) {
            unsupported = true;
            return false;
          }
        }
      }
      out.is_int = true;
      out.i = r;
      out.d = 0.0;
      return true;
    }
    const double x = cd_as_double(a), y = cd_as_double(b);
    if (x == 0.0 && y < 0.0) return false;  // ZeroDivisionError
    if (x < 0.0 && cd_finite(y) && y != floor(y)) {
      unsupported = true;  // complex result
      return false;
    }
    const double r = pow(x, y);
    if (!cd_finite(r) && cd_finite(x) && cd_finite(y)) return false;  // OverflowError
    cd_set_d(out, r);
    return true;
  }
  if (ints) {
    long long r;
    bool ok;
    if (op == '*') {
      ok = cd_mul(a.i, b.i, r);
    } else {  // |operands| <= 2^61, so + and - cannot wrap
      r = (op == '+') ? a.i + b.i : a.i - b.i;
      ok = !(r > (1LL << 61) || r < -(1LL << 61));
    }
    if (!ok) {
      unsupported = true;  // Python ints do not overflow; such magnitudes are outside the covered range
      return false;
    }
    out.is_int = true;
    out.i = r;
    out.d = 0.0;
    return true;
  }
  const double x = cd_as_double(a), y = cd_as_double(b);
  cd_set_d(out, (op == '+') ? x + y : (op == '-') ? x - y : x * y);
  return true;
}

RAGEN_HD int cd_prec(char op) {
  return (op == 'p') ? 4 : (op == 'n') ? 3 : (op == '*' || op == '/' || op == 'f' || op == '%') ? 2 : (op == '+' || op == '-') ? 1 : 0;
}

// pops one operator and applies it
RAGEN_HD bool cd_reduce(CdValue* vals, int& nv, char* ops, int& no, bool& unsupported) {
  const char op = ops[--no];
  if (op == 'n') {  // unary minus
    if (nv < 1) return false;
    if (vals[nv - 1].is_int) vals[nv - 1].i = -vals[nv - 1].i;
    else vals[nv - 1].d = -vals[nv - 1].d;
    return true;
  }
  if (nv < 2) return false;
  CdValue r;
  if (!cd_apply(op, vals[nv - 2], vals[nv - 1], r, unsupported)) return false;
  vals[nv - 2] = r;
  --nv;
  return true;
}

// A numeric literal at s[i] (a digit, or '.' followed by a digit).  Advances i past it.  Returns 0 ok, 1 SyntaxError,
// 2 outside the covered range (unsupported).
RAGEN_HD int cd_number(const uint8_t* s, int n, int& i, CdValue& out) {
  const int start = i;
  unsigned long long mant = 0;
  int sig = 0, dropped = 0;  // significant digits kept, integer digits beyond the 18th
  bool nonzero = false, is_float = false;
  while (i < n && s[i] >= '0' && s[i] <= '9') {
    if (s[i] != '0') nonzero = true;
    if (nonzero) {
      if (sig < 18) {
        mant = mant * 10 + (unsigned)(s[i] - '0');
        ++sig;
      } else {
        ++dropped;
      }
    }
    ++i;
  }
  const int int_digits = i - start;
  int frac = 0;
  if (i < n && s[i] == '.') {
    is_float = true;
    ++i;
    while (i < n && s[i] >= '0' && s[i] <= '9') {
      if (s[i] != '0') nonzero = true;
      if (nonzero && sig < 18) {
        mant = mant * 10 + (unsigned)(s[i] - '0');
        ++sig;
        ++frac;
      } else if (!nonzero) {
        ++frac;  // leading zeros of a pure fraction
      } else {
        return 2;  // more digits than the exact path carries
      }
      ++i;
    }
  }
  int e10 = 0;
  if (i < n && (s[i] == 'e' || s[i] == 'E')) {
    int j = i + 1;
    bool neg = false;
    if (j < n && (s[j] == '+' || s[j] == '-')) {
      neg = s[j] == '-';
      ++j;
    }
    if (!(j < n && s[j] >= '0' && s[j] <= '9')) return 2;  // "1e" / an identifier glued to a number
    while (j < n && s[j] >= '0' && s[j] <= '9') {
      if (e10 < 10000) e10 = e10 * 10 + (s[j] - '0');
      ++j;
    }
    if (neg) e10 = -e10;
    is_float = true;
    i = j;
  }
  if (i < n) {
    const int d = s[i];
    if (d == '.' || d == '_' || (d >= 'a' && d <= 'z') || (d >= 'A' && d <= 'Z') || d >= 0x80) return 2;  // 1_0, 1j, 0x1, 1..
  }
  if (!is_float) {
    if (int_digits > 1 && s[start] == '0' && nonzero) return 1;  // "012": SyntaxError in Python 3
    if (dropped || mant > (1ULL << 50)) return 2;
    out.is_int = true;
    out.i = (long long)mant;
    out.d = 0.0;
    return 0;
  }
  // decimal -> double on the exact path: mantissa below 2^53 times / over a power of ten up to 10^22 is correctly rounded
  if (dropped || sig > 15) return 2;
  const int p10 = e10 - frac;
  if (mant == 0) {
    cd_set_d(out, 0.0);
    return 0;
  }
  if (p10 > 22 || p10 < -22) return 2;
  double pw = 1.0;
  for (int k = 0; k < (p10 < 0 ? -p10 : p10); ++k) pw *= 10.0;  // exact up to 10^22
  cd_set_d(out, p10 >= 0 ? (double)mant * pw : (double)mant / pw);
  return 0;
}

// returns 0 = format check failed (reward 0), 1 = format ok but not correct (format_score), 2 = correct (score)
RAGEN_HD int countdown_score(const uint8_t* s, int n, const int32_t* nums, int n_nums, long long target, bool& unsupported) {
  unsupported = false;
  // ---- check_format ----
  long long found[CD_MAX_NUMS];
  int nf = 0;
  bool fmt = true;
  for (int i = 0; i < n;) {
    if (s[i] >= '0' && s[i] <= '9') {
      long long v = 0;
      bool big = false;
      while (i < n && s[i] >= '0' && s[i] <= '9') {
        if (v > (1LL << 40)) big = true;
        else v = v * 10 + (s[i] - '0');
        ++i;
      }
      if (big || nf >= CD_MAX_NUMS || nf >= n_nums) fmt = false;
      else found[nf] = v;
      ++nf;
      if (nf > CD_MAX_NUMS) nf = CD_MAX_NUMS + 1;
    } else {
      ++i;
    }
  }
  if (!fmt || nf != n_nums) return 0;
  long long want[CD_MAX_NUMS];
  for (int i = 0; i < n_nums; ++i) want[i] = nums[i];
  for (int i = 1; i < n_nums; ++i) {  // insertion sorts
    long long a = found[i], b = want[i];
    int j = i - 1, k = i - 1;
    while (j >= 0 && found[j] > a) {
      found[j + 1] = found[j];
      --j;
    }
    found[j + 1] = a;
    while (k >= 0 && want[k] > b) {
      want[k + 1] = want[k];
      --k;
    }
    want[k + 1] = b;
  }
  for (int i = 0; i < n_nums; ++i)
    if (found[i] != want[i]) return 0;
  // ---- check_correctness: operator-precedence evaluation ----
  CdValue vals[CD_STACK];
  char ops[CD_STACK];
  int nv = 0, no = 0, depth = 0;
  bool want_operand = true, any = false;
  int i = 0;
  while (i < n && (s[i] == ' ' || s[i] == '\t')) ++i;  // eval() skips leading blanks
  for (; i < n; ++i) {
    const int c = s[i];
    if (c == ' ' || c == '\t' || c == '\f') continue;
    if (c == '\n' || c == '\r') {
      if (depth > 0) continue;
      bool rest_blank = true;  // a trailing newline ends the expression
      for (int j = i; j < n; ++j)
        if (!(s[j] == ' ' || s[j] == '\t' || s[j] == '\f' || s[j] == '\n' || s[j] == '\r')) rest_blank = false;
      if (rest_blank && any && !want_operand) break;
      unsupported = true;
      return 1;
    }
    any = true;
    if (want_operand) {
      if ((c >= '0' && c <= '9') || (c == '.' && i + 1 < n && s[i + 1] >= '0' && s[i + 1] <= '9')) {
        if (nv >= CD_STACK) {
          unsupported = true;
          return 1;
        }
        const int rc = cd_number(s, n, i, vals[nv]);
        if (rc == 2) unsupported = true;
        if (rc != 0) return 1;
        --i;
        ++nv;
        want_operand = false;
      } else if (c == '+') {
        // unary plus: identity
      } else if (c == '-') {
        if (no >= CD_STACK) {
          unsupported = true;
          return 1;
        }
        ops[no++] = 'n';
      } else if (c == '(') {
        if (no >= CD_STACK) {
          unsupported = true;
          return 1;
        }
        ops[no++] = '(';
        ++depth;
      } else {
        if (!(c == ')' || c == '*' || c == '/' || c == '%')) unsupported = true;
        return 1;
      }
    } else {
      if (c == '+' || c == '-' || c == '*' || c == '/' || c == '%') {
        char op = (char)c;
        if ((c == '*' || c == '/') && i + 1 < n && s[i + 1] == c) {
          op = (c == '*') ? 'p' : 'f';  // ** or //
          ++i;
        }
        // ** is right associative: it only pops operators that bind tighter (none); the others pop their equals too
        while (no > 0 && ops[no - 1] != '(' && (op == 'p' ? cd_prec(ops[no - 1]) > cd_prec(op) : cd_prec(ops[no - 1]) >= cd_prec(op)))
          if (!cd_reduce(vals, nv, ops, no, unsupported)) return 1;
        if (no >= CD_STACK) {
          unsupported = true;
          return 1;
        }
        ops[no++] = op;
        want_operand = true;
      } else if (c == ')') {
        while (no > 0 && ops[no - 1] != '(')
          if (!cd_reduce(vals, nv, ops, no, unsupported)) return 1;
        if (no == 0) return 1;  // unmatched ')'
        --no;
        --depth;
      } else {
        // "3 4" is a SyntaxError and "3 (4)" a TypeError (int is not callable): plain exceptions
        if (!((c >= '0' && c <= '9') || c == '(' || c == '.')) unsupported = true;
        return 1;
      }
    }
  }
  if (want_operand) return 1;
  while (no > 0) {
    if (ops[no - 1] == '(') return 1;
    if (!cd_reduce(vals, nv, ops, no, unsupported)) return 1;
  }
  if (nv != 1) return 1;
  if (vals[0].is_int) return vals[0].i == target ? 2 : 1;
  const double diff = vals[0].d - (double)target;  // nan and inf compare false, as abs(result - target) < 1e-5 does
  return ((diff < 0 ? -diff : diff) < 1e-5) ? 2 : 1;
}

}  // namespace parse
