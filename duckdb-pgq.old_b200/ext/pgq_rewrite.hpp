// pgq_rewrite — the step BEFORE the hot path: turns the SQL/PGQ surface syntax
//
//     CREATE PROPERTY GRAPH g VERTEX TABLES (...) EDGE TABLES (...)
//     SELECT ... FROM GRAPH_TABLE (g MATCH p = ANY SHORTEST (a:Person)-[k:knows]->*(b:Person) COLUMNS (...)) ...
//
// into the plain SQL that calls the path's scalar functions — the query template DuckPGQ's binder produces
// (SURVEY.md §3.2 / f3): a CTE whose rows call create_csr_edge(id, |V|, CAST((SELECT sum(create_csr_vertex(...))) AS
// BIGINT), src.rowid, dst.rowid, edge.rowid), cross-joined as a one-row aggregate so that the CSR is complete before
// the first iterativelength / shortestpath / reachability call, whose vertex arguments are the ROWIDs of the vertex
// table (dense ids).  The mounted reference has no PGQ grammar (third_party/libpg_query/grammar/keywords/
// unreserved_keywords.list:126,141 holds only stock MATCH / LABEL), so this is a source-to-source rewriter that sits
// in front of Parser::ParseQuery (/root/reference/src/parser/parser.cpp:20) — INTEGRATION.md §5 shows the hook — and
// is what ext/pgq_sql.cpp runs every statement through.  No DuckDB dependency: plain C++11 on strings.
//
// Supported subset (anything else is reported as an error that names the construct, never guessed at):
//   CREATE [OR REPLACE] PROPERTY GRAPH name
//       VERTEX TABLES ( tbl [AS alias] [KEY (col)] [LABEL lbl] [NO PROPERTIES | PROPERTIES (...)] , ... )
//       EDGE TABLES   ( tbl [AS alias] [KEY (col)] SOURCE KEY (col) REFERENCES vtbl (col)
//                                       DESTINATION KEY (col) REFERENCES vtbl (col) [LABEL lbl] [...PROPERTIES...] , ... )
//   DROP PROPERTY GRAPH [IF EXISTS] name
//   GRAPH_TABLE ( name MATCH [p =] [ANY SHORTEST] (v1) e1 (v2) [e2 (v3) ...] [WHERE expr] COLUMNS ( list ) )
//       vertex  ( var [: Label] [WHERE expr] )
//       edge    -[ var [: Label] ]->   <-[ var [: Label] ]-          one hop: a join with the edge table
//               -[ var : Label ]->*  ->+  ->{l,u}  (and the <- forms)  a path: MS-BFS over that edge table's CSR
//   in COLUMNS / WHERE, for the path variable p:
//       path_length(p)              -> iterativelength(...)            (hops; with ->+ / ->{l,u} the bounded form)
//       vertices(p), element_id(p)  -> shortestpath(...)               LIST<BIGINT> of vertex rowids
//       path_with_edges(p)          -> shortestpath_with_edges(...)    [v, e, v, ...] rowids
//       cheapest_path_length(p, k.col) -> the weighted variant; the CSR is then built with the 7-argument
//                                      create_csr_edge carrying k.col (extension of this rewriter, not PGQ syntax)
//   a quantified edge without ANY SHORTEST and without a path variable only filters: reachability(...).
#pragma once

#include <algorithm>
#include <cctype>
#include <cstdlib>
#include <map>
#include <sstream>
#include <string>
#include <vector>

namespace pgq_rewrite {

struct Result {
	bool handled;                        // false: not PGQ syntax, run the statement as it is
	std::vector<std::string> statements; // statements to run instead, in order (may be empty: catalog-only statement)
	size_t main_statement;               // index of the statement whose result is the query's result
	std::string error;                   // non-empty: the statement was PGQ syntax but could not be rewritten
	Result() : handled(false), main_statement(0) {
	}
};

struct VertexTable {
	std::string table, alias, key, label;
};
struct EdgeTable {
	std::string table, alias, label;
	std::string src_key, src_table, src_ref; // SOURCE KEY (src_key) REFERENCES src_table (src_ref)
	std::string dst_key, dst_table, dst_ref;
};
struct PropertyGraph {
	std::string name;
	std::vector<VertexTable> vertices;
	std::vector<EdgeTable> edges;
};

namespace detail {

enum TokKind { T_WORD, T_NUMBER, T_STRING, T_PUNCT, T_END };
struct Tok {
	TokKind kind;
	std::string text; // words: as written; quoted identifiers keep their quotes
	std::string up;   // upper-cased (words only)
	size_t b, e;      // [b, e) in the statement
};

inline std::string upper(const std::string &s) {
	std::string r = s;
	for (size_t i = 0; i < r.size(); i++) r[i] = (char)std::toupper((unsigned char)r[i]);
	return r;
}
inline std::string lower(const std::string &s) {
	std::string r = s;
	for (size_t i = 0; i < r.size(); i++) r[i] = (char)std::tolower((unsigned char)r[i]);
	return r;
}

// SQL tokens with their source spans; comments are skipped.  Never fails: an unterminated string runs to the end.
inline std::vector<Tok> tokenize(const std::string &s) {
	std::vector<Tok> out;
	size_t i = 0, n = s.size();
	while (i < n) {
		const char c = s[i];
		if (std::isspace((unsigned char)c)) {
			i++;
			continue;
		}
		if (c == '-' && i + 1 < n && s[i + 1] == '-') {
			while (i < n && s[i] != '\n') i++;
			continue;
		}
		if (c == '/' && i + 1 < n && s[i + 1] == '*') {
			size_t j = s.find("*/", i + 2);
			i = j == std::string::npos ? n : j + 2;
			continue;
		}
		Tok t;
		t.b = i;
		if (c == '\'' || c == '"') {
			size_t j = i + 1;
			while (j < n) {
				if (s[j] == c) {
					if (j + 1 < n && s[j + 1] == c) {
						j += 2;
						continue;
					}
					break;
				}
				j++;
			}
			j = j < n ? j + 1 : n;
			t.kind = c == '\'' ? T_STRING : T_WORD;
			t.text = s.substr(i, j - i);
			t.up = t.text;
			i = j;
		} else if (std::isalpha((unsigned char)c) || c == '_') {
			size_t j = i;
			while (j < n && (std::isalnum((unsigned char)s[j]) || s[j] == '_' || s[j] == '$')) j++;
			t.kind = T_WORD;
			t.text = s.substr(i, j - i);
			t.up = upper(t.text);
			i = j;
		} else if (std::isdigit((unsigned char)c)) {
			size_t j = i;
			while (j < n && (std::isalnum((unsigned char)s[j]) || s[j] == '.')) j++;
			t.kind = T_NUMBER;
			t.text = s.substr(i, j - i);
			i = j;
		} else {
			t.kind = T_PUNCT;
			t.text = std::string(1, c);
			i++;
		}
		t.e = i;
		out.push_back(t);
	}
	Tok end;
	end.kind = T_END;
	end.b = end.e = n;
	out.push_back(end);
	return out;
}

struct ParseError {
	std::string msg;
	explicit ParseError(const std::string &m) : msg(m) {
	}
};

// Cursor over a token vector.
struct Cur {
	const std::string &src;
	const std::vector<Tok> &t;
	size_t i, end; // [i, end)
	Cur(const std::string &s, const std::vector<Tok> &toks, size_t b, size_t e) : src(s), t(toks), i(b), end(e) {
	}
	const Tok &peek(size_t k = 0) const {
		return i + k < end ? t[i + k] : t.back();
	}
	bool at_end() const {
		return i >= end || t[i].kind == T_END;
	}
	bool is_word(const char *w, size_t k = 0) const {
		const Tok &x = peek(k);
		return x.kind == T_WORD && x.up == w;
	}
	bool is_punct(char c, size_t k = 0) const {
		const Tok &x = peek(k);
		return x.kind == T_PUNCT && x.text[0] == c;
	}
	bool accept_word(const char *w) {
		if (!is_word(w)) return false;
		i++;
		return true;
	}
	bool accept_punct(char c) {
		if (!is_punct(c)) return false;
		i++;
		return true;
	}
	std::string near() const {
		if (at_end()) return "end of statement";
		return "'" + src.substr(peek().b, std::min<size_t>(24, src.size() - peek().b)) + "'";
	}
	void expect_word(const char *w) {
		if (!accept_word(w)) throw ParseError(std::string("expected ") + w + " near " + near());
	}
	void expect_punct(char c) {
		if (!accept_punct(c)) throw ParseError(std::string("expected '") + c + "' near " + near());
	}
	std::string ident(const char *what) {
		const Tok &x = peek();
		if (x.kind != T_WORD) throw ParseError(std::string("expected ") + what + " near " + near());
		i++;
		return x.text;
	}
	// index just past the ')' that closes the '(' at position `open`
	size_t match_paren(size_t open) const {
		int depth = 0;
		for (size_t k = open; k < end; k++) {
			if (t[k].kind != T_PUNCT) continue;
			const char c = t[k].text[0];
			if (c == '(') depth++;
			if (c == ')' && --depth == 0) return k + 1;
		}
		throw ParseError("unbalanced parentheses");
	}
};

inline bool same_ident(const std::string &a, const std::string &b) {
	return lower(a) == lower(b);
}

} // namespace detail
} // namespace pgq_rewrite

#include "pgq_rewrite_impl.hpp"
