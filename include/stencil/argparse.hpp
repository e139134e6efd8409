#pragma once
/* argparse::Parser — the small CLI parser every app uses.
   Surface parity with the reference's vendored thirdparty/argparse/argparse.hpp:179-365: add_flag(bool&, long,
   short)->help(), add_option(T&, long, short)->help() ("--opt value"), add_positional(T&)->required(),
   no_unrecognized(), no_consume(), need_help(), help(), parse(argc, argv) (consumes recognised args, "--" ends
   option parsing, later duplicates win). Own implementation: type-erased setters instead of a class hierarchy. */
#include <cstdint>
#include <functional>
#include <iostream>
#include <memory>
#include <sstream>
#include <string>
#include <vector>

namespace argparse {

namespace detail {
template <typename T> struct TypeName { static const char *get() { return "VALUE"; } };
template <> struct TypeName<int> { static const char *get() { return "INT"; } };
template <> struct TypeName<int64_t> { static const char *get() { return "INT"; } };
template <> struct TypeName<size_t> { static const char *get() { return "UINT"; } };
template <> struct TypeName<unsigned> { static const char *get() { return "UINT"; } };
template <> struct TypeName<float> { static const char *get() { return "FLOAT"; } };
template <> struct TypeName<double> { static const char *get() { return "FLOAT"; } };
template <> struct TypeName<std::string> { static const char *get() { return "STRING"; } };

template <typename T> inline bool convert(const std::string &s, T &out) {
  std::istringstream is(s);
  is >> out;
  return !is.fail() && is.eof();
}
template <> inline bool convert<std::string>(const std::string &s, std::string &out) {
  out = s;
  return true;
}
} // namespace detail

/* one declared argument; returned by add_* so that ->help() / ->required() chain */
class Arg {
  friend class Parser;
  std::string long_, short_, help_, type_;
  bool required_ = false;
  std::function<bool(const std::string &)> set_;

public:
  Arg *help(const std::string &h) {
    help_ = h;
    return this;
  }
  Arg *required() {
    required_ = true;
    return this;
  }
  const std::string &long_str() const { return long_; }
  const std::string &short_str() const { return short_; }
  const std::string &help_str() const { return help_; }
  bool is_required() const { return required_; }
};

class Parser {
  std::string description_, argv0_;
  bool noUnrecognized_ = false, help_ = false, consume_ = true;
  std::vector<std::unique_ptr<Arg>> flags_, opts_, posnls_;

  static Arg *find(const std::vector<std::unique_ptr<Arg>> &v, const std::string &s) {
    for (size_t i = v.size(); i-- > 0;)
      if (v[i]->long_ == s || (!v[i]->short_.empty() && v[i]->short_ == s)) return v[i].get();
    return nullptr;
  }

public:
  explicit Parser(const std::string &description = "an argparse-powered program") : description_(description) {
    add_flag(help_, "--help", "-h")->help("Print help message");
  }

  Arg *add_flag(bool &val, const std::string &l, const std::string &s = "") {
    std::unique_ptr<Arg> a(new Arg);
    a->long_ = l;
    a->short_ = s;
    bool *p = &val;
    a->set_ = [p](const std::string &) {
      *p = true;
      return true;
    };
    flags_.push_back(std::move(a));
    return flags_.back().get();
  }

  template <typename T> Arg *add_option(T &val, const std::string &l, const std::string &s = "") {
    std::unique_ptr<Arg> a(new Arg);
    a->long_ = l;
    a->short_ = s;
    a->type_ = detail::TypeName<T>::get();
    T *p = &val;
    a->set_ = [p](const std::string &str) { return detail::convert<T>(str, *p); };
    opts_.push_back(std::move(a));
    return opts_.back().get();
  }

  template <typename T> Arg *add_positional(T &val) {
    std::unique_ptr<Arg> a(new Arg);
    a->type_ = detail::TypeName<T>::get();
    T *p = &val;
    a->set_ = [p](const std::string &str) { return detail::convert<T>(str, *p); };
    posnls_.push_back(std::move(a));
    return posnls_.back().get();
  }

  void no_unrecognized() { noUnrecognized_ = true; }
  void no_consume() { consume_ = false; }
  bool need_help() const noexcept { return help_; }

  bool parse(int &argc, char **argv) {
    std::vector<char *> rest;
    if (argc > 0) {
      rest.push_back(argv[0]);
      argv0_ = argv[0] ? argv[0] : "";
    }
    size_t pi = 0;
    bool optsOkay = true;
    for (int i = 1; i < argc; ++i) {
      const std::string tok(argv[i]);
      const bool looksLikeOpt = optsOkay && tok.size() > 1 && tok[0] == '-' && !(tok[1] >= '0' && tok[1] <= '9');
      if (optsOkay && tok == "--") {
        optsOkay = false;
        continue;
      }
      if (looksLikeOpt) {
        if (Arg *o = find(opts_, tok)) {
          if (i + 1 >= argc) {
            std::cerr << "missing value for " << tok << "\n";
            return false;
          }
          if (!o->set_(argv[++i])) {
            std::cerr << "bad value '" << argv[i] << "' for " << tok << "\n";
            return false;
          }
          continue;
        }
        if (Arg *f = find(flags_, tok)) {
          f->set_("");
          continue;
        }
        rest.push_back(argv[i]);
        if (noUnrecognized_) {
          std::cerr << "unrecognized " << tok << "\n";
          return false;
        }
      } else if (pi < posnls_.size()) {
        if (!posnls_[pi]->set_(tok)) {
          std::cerr << "bad value '" << tok << "' for positional argument " << pi << "\n";
          return false;
        }
        ++pi;
      } else {
        rest.push_back(argv[i]);
        std::cerr << "encountered unexpected positional argument " << pi << ": " << tok << "\n";
      }
    }
    if (!help_)
      for (; pi < posnls_.size(); ++pi)
        if (posnls_[pi]->required_) {
          std::cerr << "missing required positional argument " << pi << "\n";
          return false;
        }
    if (consume_) {
      argc = int(rest.size());
      for (int i = 0; i < argc; ++i) argv[i] = rest[size_t(i)];
    }
    return true;
  }

  std::string help() const {
    std::ostringstream ss;
    ss << "Usage: " << argv0_ << " [OPTION]...";
    for (const auto &p : posnls_) ss << " " << (p->required_ ? "" : "[") << p->type_ << (p->required_ ? "" : "]");
    ss << "\n" << description_ << "\n";
    for (const auto &o : opts_) {
      ss << "  ";
      if (!o->short_.empty()) ss << o->short_ << ", ";
      ss << o->long_ << "\t" << o->type_ << "\t\t" << o->help_ << "\n";
    }
    for (const auto &f : flags_) {
      ss << "  ";
      if (!f->short_.empty()) ss << f->short_ << ", ";
      ss << f->long_ << "\t\t" << f->help_ << "\n";
    }
    return ss.str();
  }
};

} // namespace argparse
