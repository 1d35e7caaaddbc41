// caffe_text.hpp — protobuf TEXT format (prototxt) -> protobuf wire format for the Caffe messages on this path.
// Backs parse_net_prototxt / parse_solver_prototxt (reference: libccaffe/ccaffe.cpp:275-304, which calls
// caffe::ReadProtoFromTextFile, caffe/src/caffe/util/io.cpp:34-43, and re-serialises the message for the JVM).
// No protobuf library exists in this build, so the parser walks generated field tables (caffe_schema.inc, produced by
// tools/gen_schema_inc.py from the Python schema that tests pin against the reference's own descriptor).
// Grammar handled: `name: scalar`, `name { ... }` / `name: { ... }` / `name < ... >`, `name: [a, b]`, `#` comments,
// `,` / `;` separators, single- and double-quoted strings with C escapes, adjacent string literals, enum names or
// numbers, hex / octal integers, floats with an `f` suffix, inf / nan.  An unknown field is an error, as in protobuf.
#pragma once
#include <stdint.h>

#include <cmath>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "caffe_proto.hpp"

namespace cptxt {
namespace cp = caffe_proto;

enum TxtType { T_INT32, T_INT64, T_UINT32, T_BOOL, T_FLOAT, T_DOUBLE, T_STRING, T_ENUM, T_MESSAGE };
struct TxtEnumValue { const char* name; int value; };
struct TxtEnum { const char* name; const TxtEnumValue* values; int n; };
struct TxtField { const char* name; int number; TxtType type; int ref; int repeated; };
struct TxtMessage { const char* name; const TxtField* fields; int n; };

#include "caffe_schema.inc"

struct Token { enum Kind { IDENT, STRING, PUNCT, END } kind; std::string text; size_t line; };

struct Lexer {
  const std::string& s;
  size_t i = 0, line = 1;
  explicit Lexer(const std::string& src) : s(src) {}
  [[noreturn]] void fail(const std::string& what) const {
    throw std::runtime_error("prototxt line " + std::to_string(line) + ": " + what);
  }
  Token next() {
    for (;;) {                                            // white space and comments
      while (i < s.size() && (s[i] == ' ' || s[i] == '\t' || s[i] == '\r' || s[i] == '\n')) { if (s[i] == '\n') ++line; ++i; }
      if (i < s.size() && s[i] == '#') { while (i < s.size() && s[i] != '\n') ++i; continue; }
      break;
    }
    if (i >= s.size()) return {Token::END, "", line};
    const char c = s[i];
    if (strchr("{}:<>[],;", c)) { ++i; return {Token::PUNCT, std::string(1, c), line}; }
    if (c == '"' || c == '\'') {
      std::string out;
      ++i;
      while (i < s.size() && s[i] != c) {
        if (s[i] == '\n') fail("newline inside a string literal");
        if (s[i] != '\\') { out.push_back(s[i++]); continue; }
        if (++i >= s.size()) break;
        const char e = s[i++];
        switch (e) {
          case 'n': out.push_back('\n'); break;
          case 't': out.push_back('\t'); break;
          case 'r': out.push_back('\r'); break;
          case '0': case '1': case '2': case '3': case '4': case '5': case '6': case '7': {
            int v = e - '0', k = 1;
            while (k < 3 && i < s.size() && s[i] >= '0' && s[i] <= '7') { v = v * 8 + (s[i++] - '0'); ++k; }
            out.push_back((char)v);
            break;
          }
          case 'x': {
            int v = 0, k = 0;
            while (k < 2 && i < s.size() && isxdigit((unsigned char)s[i])) { v = v * 16 + (isdigit((unsigned char)s[i]) ? s[i] - '0' : (tolower(s[i]) - 'a' + 10)); ++i; ++k; }
            out.push_back((char)v);
            break;
          }
          default: out.push_back(e);                      // \\ \" \' and anything else: the character itself
        }
      }
      if (i >= s.size()) fail("unterminated string literal");
      ++i;
      return {Token::STRING, out, line};
    }
    const size_t b = i;
    while (i < s.size() && !strchr(" \t\r\n{}:<>[],;\"'#", s[i])) ++i;
    return {Token::IDENT, s.substr(b, i - b), line};
  }
};

struct Parser {
  Lexer lex;
  Token tok;
  explicit Parser(const std::string& src) : lex(src) { tok = lex.next(); }
  void advance() { tok = lex.next(); }
  bool is_punct(char c) const { return tok.kind == Token::PUNCT && tok.text[0] == c; }

  static const TxtField* find(const TxtMessage& m, const std::string& name) {
    for (int i = 0; i < m.n; ++i) if (name == m.fields[i].name) return &m.fields[i];
    return nullptr;
  }
  void scalar(const TxtField& f, cp::Writer& w) {
    if (f.type == T_STRING) {
      if (tok.kind != Token::STRING) lex.fail(std::string("field ") + f.name + " expects a string");
      std::string v = tok.text;
      advance();
      while (tok.kind == Token::STRING) { v += tok.text; advance(); }      // adjacent literals concatenate
      w.bytes(f.number, v);
      return;
    }
    if (tok.kind != Token::IDENT) lex.fail(std::string("field ") + f.name + " expects a value");
    const std::string t = tok.text;
    advance();
    switch (f.type) {
      case T_BOOL:
        if (t == "true" || t == "True" || t == "t" || t == "1") w.i32(f.number, 1);
        else if (t == "false" || t == "False" || t == "f" || t == "0") w.i32(f.number, 0);
        else lex.fail("bad bool " + t);
        break;
      case T_ENUM: {
        const TxtEnum& e = kTxtEnums[f.ref];
        for (int i = 0; i < e.n; ++i) if (t == e.values[i].name) { w.i32(f.number, e.values[i].value); return; }
        char* end = nullptr;
        const long v = strtol(t.c_str(), &end, 0);
        if (!end || *end) lex.fail("unknown value " + t + " of enum " + e.name);
        w.i32(f.number, v);
        break;
      }
      case T_FLOAT: case T_DOUBLE: {
        std::string u = t;
        if (u.size() > 1 && (u.back() == 'f' || u.back() == 'F') && u != "inf" && u != "-inf") u.pop_back();
        char* end = nullptr;
        const double v = strtod(u.c_str(), &end);
        if (!end || *end) lex.fail("bad number " + t);
        if (f.type == T_FLOAT) w.f32(f.number, (float)v);
        else { w.tag(f.number, 1); char b[8]; memcpy(b, &v, 8); w.out.append(b, 8); }
        break;
      }
      default: {                                           // int32 / int64 / uint32: decimal, hex, octal
        char* end = nullptr;
        const long long v = strtoll(t.c_str(), &end, 0);
        if (!end || *end) lex.fail("bad integer " + t);
        w.i32(f.number, (int64_t)v);                       // negative values: 10-byte two's-complement varint, as protobuf
      }
    }
  }
  // fields of message `m` until `closer` (0 = end of input)
  std::string message(const TxtMessage& m, char closer) {
    cp::Writer w;
    for (;;) {
      if (tok.kind == Token::END) { if (closer) lex.fail(std::string("unexpected end of input inside ") + m.name); return w.out; }
      if (closer && is_punct(closer)) { advance(); return w.out; }
      if (is_punct(',') || is_punct(';')) { advance(); continue; }
      if (tok.kind != Token::IDENT) lex.fail("expected a field name, got '" + tok.text + "'");
      const TxtField* f = find(m, tok.text);
      if (!f) lex.fail(std::string("message ") + m.name + " has no field named '" + tok.text + "' on the supported path");
      advance();
      const bool colon = is_punct(':');
      if (colon) advance();
      if (f->type == T_MESSAGE) {
        if (!is_punct('{') && !is_punct('<')) lex.fail(std::string("expected '{' after ") + f->name);
        const char close = is_punct('{') ? '}' : '>';
        advance();
        w.bytes(f->number, message(kTxtMessages[f->ref], close));
      } else {
        if (!colon) lex.fail(std::string("expected ':' after ") + f->name);
        if (is_punct('[')) {
          advance();
          while (!is_punct(']')) {
            if (tok.kind == Token::END) lex.fail("unterminated list");
            if (is_punct(',')) { advance(); continue; }
            scalar(*f, w);
          }
          advance();
        } else scalar(*f, w);
      }
    }
  }
};

// text of a `root` message ("NetParameter" / "SolverParameter") -> its wire encoding; throws std::runtime_error
inline std::string text_to_wire(const char* root, const std::string& text) {
  for (int i = 0; i < kTxtNumMessages; ++i)
    if (!strcmp(kTxtMessages[i].name, root)) {
      Parser p(text);
      return p.message(kTxtMessages[i], 0);
    }
  throw std::runtime_error(std::string("prototxt: unknown root message ") + root);
}

}  // namespace cptxt
