// Byte-level logic of the response parser and the Countdown scorer, as pure host/device functions.
// The kernels in parse.cu call these; tests/test_hostsim.py runs the same code on the host against the oracle.
//
// Follows ContextManager._parse_response (ragen/llm_agent/ctx_manager.py:148-173), EnvStateManager.
// _extract_map_valid_actions (ragen/llm_agent/es_manager.py:230-240) and CountdownEnv.compute_reward /
// check_format / check_correctness (ragen/env/countdown/env.py:9-21, 69-78).  Text is UTF-8.
#pragma once
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
// expressions: decimal integers (leading zeros only for zero, as Python 3), binary + - * / (true division), unary
// + -, parentheses, blanks (space, tab, form feed; newlines inside parentheses or trailing).  Anything else Python
// might accept (** // % floats names comparisons ...) is reported through `unsupported` and scored as an exception.
constexpr int CD_MAX_NUMS = 8;
constexpr int CD_STACK = 32;

struct CdValue {
  bool is_int;
  long long i;
  double d;
};
RAGEN_HD double cd_as_double(const CdValue& v) { return v.is_int ? (double)v.i : v.d; }

// applies a binary operator; false on ZeroDivisionError / int64 overflow (overflow also sets unsupported)
RAGEN_HD bool cd_apply(char op, const CdValue& a, const CdValue& b, CdValue& out, bool& unsupported) {
  if (op == '/') {
    const double den = cd_as_double(b);
    if (den == 0.0) return false;
    out.is_int = false;
    out.d = cd_as_double(a) / den;
    out.i = 0;
    return true;
  }
  if (a.is_int && b.is_int) {
    long long r;
    bool ovf;
#if defined(__CUDA_ARCH__)
    // |operands| are checked to stay below 2^62 so + and - cannot wrap; * is checked by division
    if (op == '*') {
      r = a.i * b.i;
      ovf = (a.i != 0 && (r / a.i != b.i || (a.i == -1 && b.i == (-9223372036854775807LL - 1))));
    } else {
      r = (op == '+') ? a.i + b.i : a.i - b.i;
      ovf = false;
    }
#else
    ovf = (op == '*') ? __builtin_mul_overflow(a.i, b.i, &r) : (op == '+') ? __builtin_add_overflow(a.i, b.i, &r)
                                                                           : __builtin_sub_overflow(a.i, b.i, &r);
#endif
    if (ovf || r > (1LL << 61) || r < -(1LL << 61)) {
      unsupported = true;  // Python ints do not overflow; such magnitudes are outside the covered range
      return false;
    }
    out.is_int = true;
    out.i = r;
    out.d = 0.0;
    return true;
  }
  const double x = cd_as_double(a), y = cd_as_double(b);
  out.is_int = false;
  out.i = 0;
  out.d = (op == '+') ? x + y : (op == '-') ? x - y : x * y;
  return true;
}

RAGEN_HD int cd_prec(char op) { return (op == 'n') ? 3 : (op == '*' || op == '/') ? 2 : (op == '+' || op == '-') ? 1 : 0; }

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
      if (c >= '0' && c <= '9') {
        const int start = i;
        long long v = 0;
        bool nonzero = false;
        while (i < n && s[i] >= '0' && s[i] <= '9') {
          if (s[i] != '0') nonzero = true;
          if (v > (1LL << 50)) {
            unsupported = true;
            return 1;
          }
          v = v * 10 + (s[i] - '0');
          ++i;
        }
        if (i - start > 1 && s[start] == '0' && nonzero) return 1;  // "012": SyntaxError in Python 3
        if (i < n) {
          const int d = s[i];
          if (d == '.' || d == '_' || (d >= 'a' && d <= 'z') || (d >= 'A' && d <= 'Z') || d >= 0x80) {
            unsupported = true;  // float / complex / digit separators / identifiers glued to a number
            return 1;
          }
        }
        --i;
        if (nv >= CD_STACK) {
          unsupported = true;
          return 1;
        }
        vals[nv].is_int = true;
        vals[nv].i = v;
        vals[nv].d = 0.0;
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
        if (!(c == ')' || c == '*' || c == '/')) unsupported = true;
        return 1;
      }
    } else {
      if (c == '+' || c == '-' || c == '*' || c == '/') {
        if ((c == '*' || c == '/') && i + 1 < n && s[i + 1] == c) {
          unsupported = true;  // ** or //
          return 1;
        }
        while (no > 0 && ops[no - 1] != '(' && cd_prec(ops[no - 1]) >= cd_prec((char)c))
          if (!cd_reduce(vals, nv, ops, no, unsupported)) return 1;
        if (no >= CD_STACK) {
          unsupported = true;
          return 1;
        }
        ops[no++] = (char)c;
        want_operand = true;
      } else if (c == ')') {
        while (no > 0 && ops[no - 1] != '(')
          if (!cd_reduce(vals, nv, ops, no, unsupported)) return 1;
        if (no == 0) return 1;  // unmatched ')'
        --no;
        --depth;
      } else {
        // "3 4" is a SyntaxError and "3 (4)" a TypeError (int is not callable): plain exceptions
        if (!((c >= '0' && c <= '9') || c == '(')) unsupported = true;
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
  const double diff = vals[0].d - (double)target;
  return ((diff < 0 ? -diff : diff) < 1e-5) ? 2 : 1;
}

}  // namespace parse
