// predicate.cpp — evaluates the pushed-down WHERE of the optimizer rewrite into the prefilter bitmap.
// Input grammar = exactly what ExpressionToLancePredicate emits (src/lance_optimizer.cpp:204-344):
//   col <op> const | const <op> col   (=, !=, <, >, <=, >=)      children of AND;  OR children in parentheses
//   col IS [NOT] NULL | NOT (<expr>) | col [NOT] IN (c1, c2, ..) | BETWEEN as "col >= a AND col <= b"
// Constants as FormatConstant writes them (:138-180): 'text' with '' escapes, integers, floats, true/false, NULL.
// Semantics = SQL three-valued logic, a row passes only when the predicate is TRUE (prefilter semantics pinned by
// test/sql/lance_optimizer_filter.test:23-99).
#include <cctype>
#include <cmath>
#include <cstdlib>
#include <cstring>

#include "lance_host.hpp"

namespace lance_host {
namespace {

typedef std::vector<int8_t> Tri; // 1 true, 0 false, -1 null

struct Literal {
	enum Kind { NUL, STR, INT, FLT, BOOL } kind = NUL;
	std::string s;
	int64_t i = 0;
	double d = 0;
};

struct Operand {
	bool is_col = false;
	std::string col;
	Literal lit;
};

struct Parser {
	const std::string &src;
	size_t pos = 0;
	const ColumnSet &cols;
	int64_t n;
	Parser(const std::string &s, const ColumnSet &c, int64_t n_) : src(s), cols(c), n(n_) {}

	[[noreturn]] void fail(const std::string &why) const {
		throw BinderException("lance predicate: " + why + " at offset " + std::to_string(pos) + " in \"" + src + "\"");
	}
	void skip() {
		while (pos < src.size() && isspace((unsigned char)src[pos])) pos++;
	}
	bool eof() {
		skip();
		return pos >= src.size();
	}
	bool peek_kw(const char *kw) {
		skip();
		size_t len = strlen(kw);
		if (pos + len > src.size()) return false;
		for (size_t i = 0; i < len; i++)
			if (toupper((unsigned char)src[pos + i]) != kw[i]) return false;
		if (pos + len < src.size() && (isalnum((unsigned char)src[pos + len]) || src[pos + len] == '_')) return false;
		return true;
	}
	bool take_kw(const char *kw) {
		if (!peek_kw(kw)) return false;
		pos += strlen(kw);
		return true;
	}
	bool take_ch(char c) {
		skip();
		if (pos < src.size() && src[pos] == c) {
			pos++;
			return true;
		}
		return false;
	}

	Operand operand() {
		skip();
		Operand o;
		if (pos >= src.size()) fail("operand expected");
		char c = src[pos];
		if (c == '\'') {
			pos++;
			o.lit.kind = Literal::STR;
			while (true) {
				if (pos >= src.size()) fail("unterminated string");
				if (src[pos] == '\'') {
					if (pos + 1 < src.size() && src[pos + 1] == '\'') {
						o.lit.s += '\'';
						pos += 2;
						continue;
					}
					pos++;
					break;
				}
				o.lit.s += src[pos++];
			}
			return o;
		}
		if (isdigit((unsigned char)c) || c == '-' || c == '+' || c == '.') {
			size_t b = pos;
			if (c == '-' || c == '+') pos++;
			bool flt = false;
			while (pos < src.size() && (isdigit((unsigned char)src[pos]) || src[pos] == '.' || src[pos] == 'e' || src[pos] == 'E' ||
			                            ((src[pos] == '-' || src[pos] == '+') && (src[pos - 1] == 'e' || src[pos - 1] == 'E')))) {
				if (!isdigit((unsigned char)src[pos])) flt = true;
				pos++;
			}
			std::string tok = src.substr(b, pos - b);
			if (flt) {
				o.lit.kind = Literal::FLT;
				o.lit.d = strtod(tok.c_str(), nullptr);
			} else {
				o.lit.kind = Literal::INT;
				o.lit.i = strtoll(tok.c_str(), nullptr, 10);
				o.lit.d = (double)o.lit.i;
			}
			return o;
		}
		if (take_kw("TRUE")) {
			o.lit.kind = Literal::BOOL, o.lit.i = 1;
			return o;
		}
		if (take_kw("FALSE")) {
			o.lit.kind = Literal::BOOL, o.lit.i = 0;
			return o;
		}
		if (take_kw("NULL")) return o;
		if (isalpha((unsigned char)c) || c == '_' || c == '"') {
			o.is_col = true;
			if (c == '"') {
				pos++;
				while (pos < src.size() && src[pos] != '"') o.col += src[pos++];
				if (pos >= src.size()) fail("unterminated identifier");
				pos++;
			} else {
				while (pos < src.size() && (isalnum((unsigned char)src[pos]) || src[pos] == '_')) o.col += src[pos++];
			}
			if (!cols.count(o.col)) fail("unknown column \"" + o.col + "\"");
			return o;
		}
		fail("unexpected character");
	}

	// -1 / 0 / 1 comparison of row r of column c with a literal; returns false when either side is NULL
	bool compare(const Column &c, int64_t r, const Literal &l, int &out) const {
		if (l.kind == Literal::NUL || r >= (int64_t)c.size() || !c.valid[r]) return false;
		switch (c.type) {
		case ColumnType::VARCHAR: {
			if (l.kind != Literal::STR) return false;
			int v = c.s[r].compare(l.s);
			out = v < 0 ? -1 : (v > 0 ? 1 : 0);
			return true;
		}
		case ColumnType::BIGINT:
			if (l.kind == Literal::INT || l.kind == Literal::BOOL) {
				out = c.i[r] < l.i ? -1 : (c.i[r] > l.i ? 1 : 0);
				return true;
			}
			if (l.kind == Literal::FLT) {
				double v = (double)c.i[r];
				out = v < l.d ? -1 : (v > l.d ? 1 : 0);
				return true;
			}
			return false;
		case ColumnType::DOUBLE: {
			if (l.kind != Literal::INT && l.kind != Literal::FLT) return false;
			double v = c.d[r];
			out = v < l.d ? -1 : (v > l.d ? 1 : 0);
			return true;
		}
		case ColumnType::BOOLEAN:
			if (l.kind != Literal::BOOL && l.kind != Literal::INT) return false;
			out = c.i[r] < l.i ? -1 : (c.i[r] > l.i ? 1 : 0);
			return true;
		}
		return false;
	}

	static bool apply(const std::string &op, int cmp) {
		if (op == "=") return cmp == 0;
		if (op == "!=" || op == "<>") return cmp != 0;
		if (op == "<") return cmp < 0;
		if (op == ">") return cmp > 0;
		if (op == "<=") return cmp <= 0;
		return cmp >= 0;
	}
	static std::string flip(const std::string &op) {
		if (op == "<") return ">";
		if (op == ">") return "<";
		if (op == "<=") return ">=";
		if (op == ">=") return "<=";
		return op;
	}

	Tri primary() {
		if (take_ch('(')) {
			Tri t = expr();
			if (!take_ch(')')) fail("')' expected");
			return t;
		}
		Operand a = operand();
		Tri out((size_t)n, (int8_t)-1);
		if (a.is_col && peek_kw("IS")) {
			take_kw("IS");
			bool neg = take_kw("NOT");
			if (!take_kw("NULL")) fail("NULL expected after IS");
			const Column &c = cols.at(a.col);
			for (int64_t r = 0; r < n; r++) {
				bool isnull = r >= (int64_t)c.size() || !c.valid[r];
				out[r] = (isnull != neg) ? 1 : 0;
			}
			return out;
		}
		if (a.is_col && (peek_kw("IN") || peek_kw("NOT"))) {
			bool neg = take_kw("NOT");
			if (!take_kw("IN")) fail("IN expected");
			if (!take_ch('(')) fail("'(' expected after IN");
			std::vector<Literal> lits;
			do {
				Operand o = operand();
				if (o.is_col) fail("constant expected in IN list");
				lits.push_back(o.lit);
			} while (take_ch(','));
			if (!take_ch(')')) fail("')' expected");
			const Column &c = cols.at(a.col);
			for (int64_t r = 0; r < n; r++) {
				if (r >= (int64_t)c.size() || !c.valid[r]) continue; // NULL IN (..) is NULL
				bool hit = false, saw_null = false;
				for (auto &l : lits) {
					int cmp;
					if (l.kind == Literal::NUL) saw_null = true;
					else if (compare(c, r, l, cmp) && cmp == 0) hit = true;
				}
				if (hit) out[r] = neg ? 0 : 1;
				else if (saw_null) out[r] = -1;
				else out[r] = neg ? 1 : 0;
			}
			return out;
		}
		// comparison
		skip();
		std::string op;
		static const char *ops[] = {"<=", ">=", "!=", "<>", "=", "<", ">"};
		for (const char *o : ops)
			if (src.compare(pos, strlen(o), o) == 0) {
				op = o;
				pos += strlen(o);
				break;
			}
		if (op.empty()) fail("comparison operator expected");
		Operand b = operand();
		if (a.is_col == b.is_col) fail("a comparison needs one column and one constant"); // :216-232
		const Column &c = cols.at(a.is_col ? a.col : b.col);
		const Literal &l = a.is_col ? b.lit : a.lit;
		const std::string eff = a.is_col ? op : flip(op);
		for (int64_t r = 0; r < n; r++) {
			int cmp;
			if (compare(c, r, l, cmp)) out[r] = apply(eff, cmp) ? 1 : 0;
		}
		return out;
	}
	Tri not_expr() {
		if (take_kw("NOT")) {
			Tri t = not_expr();
			for (auto &v : t)
				if (v >= 0) v = (int8_t)(1 - v);
			return t;
		}
		return primary();
	}
	Tri and_expr() {
		Tri t = not_expr();
		while (take_kw("AND")) {
			Tri u = not_expr();
			for (int64_t r = 0; r < n; r++) {
				if (t[r] == 0 || u[r] == 0) t[r] = 0;
				else if (t[r] < 0 || u[r] < 0) t[r] = -1;
				else t[r] = 1;
			}
		}
		return t;
	}
	Tri expr() {
		Tri t = and_expr();
		while (take_kw("OR")) {
			Tri u = and_expr();
			for (int64_t r = 0; r < n; r++) {
				if (t[r] == 1 || u[r] == 1) t[r] = 1;
				else if (t[r] < 0 || u[r] < 0) t[r] = -1;
				else t[r] = 0;
			}
		}
		return t;
	}
};

} // namespace

std::vector<uint32_t> EvaluatePredicate(const std::string &predicate, const ColumnSet &cols, int64_t n) {
	Parser p(predicate, cols, n);
	Tri t = p.expr();
	if (!p.eof()) p.fail("trailing input");
	std::vector<uint32_t> bm((size_t)((n + 31) / 32) + 1, 0u);
	for (int64_t r = 0; r < n; r++)
		if (t[r] == 1) bm[(size_t)(r >> 5)] |= 1u << (r & 31);
	return bm;
}

} // namespace lance_host
