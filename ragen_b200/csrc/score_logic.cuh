// Answer scoring of the text environments, host/device logic (no CUDA dependency beyond RAGEN_HD).
//
// Restates MetaMathQAEnv._check_answer (ragen/env/metamathqa/env.py:55-66) and the Static env's scorers
// compute_score_exact_match / compute_score_numeric / compute_score_multiple_choice with normalize_text
// (ragen/env/static/utils.py:7-12, 69-132) on UTF-8 byte strings.
//
// Exact for ASCII input.  Python lower-cases, upper-cases and classifies whitespace / digits by Unicode tables; here
// bytes >= 0x80 are kept as they are, and the `unsupported` output bit is set whenever either string holds such a byte
// (or a number with more than 15 significant digits, see numeric below), so a caller can tell the exact results apart.
#pragma once
#include <stdint.h>

#include "parse_logic.cuh"  // RAGEN_HD, parse::space_at

namespace score {

enum Mode { METAMATH = 0, EXACT = 1, NUMERIC = 2, CHOICE = 3 };
enum Bits { CORRECT = 1, VALID = 2, TEXT_MATCH = 4, UNSUPPORTED = 128 };

RAGEN_HD int lower(int c) { return (c >= 'A' && c <= 'Z') ? c + 32 : c; }
// string.punctuation: the 32 ASCII punctuation characters
RAGEN_HD bool is_punct(int c) { return (c >= 33 && c <= 47) || (c >= 58 && c <= 64) || (c >= 91 && c <= 96) || (c >= 123 && c <= 126); }
RAGEN_HD bool has_high(const uint8_t* s, int n) {
  for (int i = 0; i < n; ++i)
    if (s[i] >= 0x80) return true;
  return false;
}

// next kept byte of s at or after i (lower-cased), or -1 at the end; whitespace (re \s) is always dropped, punctuation
// when `drop_punct`
RAGEN_HD int next_kept(const uint8_t* s, int n, int& i, bool drop_punct) {
  while (i < n) {
    const int k = parse::space_at(s + i, s + n);
    if (k) {
      i += k;
      continue;
    }
    const int c = s[i];
    if (drop_punct && is_punct(c)) {
      ++i;
      continue;
    }
    ++i;
    return lower(c);
  }
  return -1;
}

// normalised(a) == normalised(b); *a_nonempty = normalised(a) != ""
RAGEN_HD bool norm_equal(const uint8_t* a, int na, const uint8_t* b, int nb, bool drop_punct, bool* a_nonempty) {
  int i = 0, j = 0;
  bool first = true, equal = true;
  for (;;) {
    const int x = next_kept(a, na, i, drop_punct);
    if (first) {
      *a_nonempty = x >= 0;
      first = false;
    }
    const int y = next_kept(b, nb, j, drop_punct);
    if (x != y) equal = false;
    if (x < 0 || y < 0) break;
    if (!equal) break;
  }
  return equal;
}

// re.search(r'(\d+(?:\.\d+)?)', s): first ASCII digit run with an optional fraction.  Returns false when there is none.
struct Number {
  int int_s, int_e, frac_s, frac_e;  // [int_s, int_e) digits, [frac_s, frac_e) fraction digits (empty when absent)
};
RAGEN_HD bool is_digit(int c) { return c >= '0' && c <= '9'; }
RAGEN_HD bool first_number(const uint8_t* s, int n, Number& out) {
  int i = 0;
  while (i < n && !is_digit(s[i])) ++i;
  if (i >= n) return false;
  out.int_s = i;
  while (i < n && is_digit(s[i])) ++i;
  out.int_e = i;
  out.frac_s = out.frac_e = i;
  if (i + 1 < n && s[i] == '.' && is_digit(s[i + 1])) {
    out.frac_s = i + 1;
    int j = i + 1;
    while (j < n && is_digit(s[j])) ++j;
    out.frac_e = j;
  }
  return true;
}
// float(x) == float(y) decided on the decimal strings: leading zeros of the integer part and trailing zeros of the fraction
// do not change the value; two canonical decimals of at most 15 significant digits are equal as doubles iff they are equal as
// strings.  `wide` is set when either has more, where two different decimals may round to the same double.
RAGEN_HD bool decimal_equal(const uint8_t* a, Number x, const uint8_t* b, Number y, bool* wide) {
  while (x.int_s < x.int_e && a[x.int_s] == '0') ++x.int_s;
  while (y.int_s < y.int_e && b[y.int_s] == '0') ++y.int_s;
  while (x.frac_e > x.frac_s && a[x.frac_e - 1] == '0') --x.frac_e;
  while (y.frac_e > y.frac_s && b[y.frac_e - 1] == '0') --y.frac_e;
  const int xi = x.int_e - x.int_s, yi = y.int_e - y.int_s, xf = x.frac_e - x.frac_s, yf = y.frac_e - y.frac_s;
  if (xi + xf > 15 || yi + yf > 15) *wide = true;
  if (xi != yi || xf != yf) return false;
  for (int k = 0; k < xi; ++k)
    if (a[x.int_s + k] != b[y.int_s + k]) return false;
  for (int k = 0; k < xf; ++k)
    if (a[x.frac_s + k] != b[y.frac_s + k]) return false;
  return true;
}

// re.search(r'([A-D])', s.upper())
RAGEN_HD int first_choice(const uint8_t* s, int n) {
  for (int i = 0; i < n; ++i) {
    const int c = s[i];
    if (c >= 'A' && c <= 'D') return c;
    if (c >= 'a' && c <= 'd') return c - 32;
  }
  return 0;
}

// -> Bits
RAGEN_HD int score_pair(int mode, const uint8_t* pred, int np, const uint8_t* label, int nl) {
  int out = (has_high(pred, np) || has_high(label, nl)) ? UNSUPPORTED : 0;
  bool nonempty = false;
  if (mode == METAMATH) {
    // user_answer.strip() then all whitespace removed and lower(): the strip is subsumed.  `if self.correct_answer:` -
    // an empty label is never correct (:60-64); is_valid = normalised answer != "" (:65)
    const bool eq = norm_equal(pred, np, label, nl, false, &nonempty);
    if (eq && nl > 0) out |= CORRECT;
    if (nonempty) out |= VALID;
    return out;
  }
  const bool text = norm_equal(pred, np, label, nl, true, &nonempty);
  if (text) out |= TEXT_MATCH;
  if (mode == EXACT) {
    if (text) out |= CORRECT;
    if (nonempty) out |= VALID;  // len(norm_pred) > 0 (:75)
    return out;
  }
  if (mode == NUMERIC) {
    Number x, y;
    const bool px = first_number(pred, np, x), py = first_number(label, nl, y);
    bool wide = false;
    bool correct = px && py && decimal_equal(pred, x, label, y, &wide);
    if (wide) out |= UNSUPPORTED;
    if (correct || text) out |= CORRECT;  // is_correct or text_match (:104-106)
    if (px) out |= VALID;
    return out;
  }
  const int cp = first_choice(pred, np), cl = first_choice(label, nl);
  if (cp && cl ? cp == cl : text) out |= CORRECT;  // fallback to text comparison when either has no letter (:123-128)
  if (cp) out |= VALID;
  return out;
}

}  // namespace score
