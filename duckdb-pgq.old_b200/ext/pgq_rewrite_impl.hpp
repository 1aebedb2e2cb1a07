// Implementation of pgq_rewrite.hpp (included by it; see there for the contract and the supported subset).
#pragma once

namespace pgq_rewrite {
namespace detail {

struct PVertex {
	std::string var;
	const VertexTable *tbl;
	std::string where; // raw SQL text, may be empty
};
struct PEdge {
	std::string var;
	const EdgeTable *tbl;
	bool reversed;   // <-[ ]-
	bool quantified; // * + {l,u}: a path, answered by MS-BFS over the edge table's CSR
	long lower, upper; // upper < 0: unbounded
	int csr_id;
	std::string weight_col; // set when cheapest_path_length(p, var.col) is used
	bool weight_double;
	PEdge() : tbl(NULL), reversed(false), quantified(false), lower(0), upper(-1), csr_id(-1), weight_double(false) {
	}
};

inline const VertexTable *find_vertex(const PropertyGraph &g, const std::string &label_or_table) {
	for (size_t i = 0; i < g.vertices.size(); i++)
		if (same_ident(g.vertices[i].label, label_or_table)) return &g.vertices[i];
	for (size_t i = 0; i < g.vertices.size(); i++)
		if (same_ident(g.vertices[i].table, label_or_table) || same_ident(g.vertices[i].alias, label_or_table))
			return &g.vertices[i];
	return NULL;
}
inline const EdgeTable *find_edge(const PropertyGraph &g, const std::string &label_or_table) {
	for (size_t i = 0; i < g.edges.size(); i++)
		if (same_ident(g.edges[i].label, label_or_table)) return &g.edges[i];
	for (size_t i = 0; i < g.edges.size(); i++)
		if (same_ident(g.edges[i].table, label_or_table) || same_ident(g.edges[i].alias, label_or_table))
			return &g.edges[i];
	return NULL;
}

// raw source text of tokens [b, e)
inline std::string span_text(const std::string &src, const std::vector<Tok> &t, size_t b, size_t e) {
	if (b >= e) return std::string();
	return src.substr(t[b].b, t[e - 1].e - t[b].b);
}

inline std::string num(long x) {
	std::ostringstream o;
	o << x;
	return o.str();
}

} // namespace detail

class Rewriter {
public:
	Rewriter() : next_csr_(1000) {
	}
	// csr ids handed to the generated statements start here (the ids a user passes by hand stay free)
	void set_first_csr_id(int id) {
		next_csr_ = id;
	}
	const std::map<std::string, PropertyGraph> &graphs() const {
		return graphs_;
	}

	Result rewrite(const std::string &sql) {
		using namespace detail;
		Result r;
		const std::vector<Tok> toks = tokenize(sql);
		Cur c(sql, toks, 0, toks.size());
		try {
			if (c.is_word("CREATE")) {
				size_t k = 1;
				if (c.is_word("OR", 1) && c.is_word("REPLACE", 2)) k = 3;
				if (c.is_word("PROPERTY", k) && c.is_word("GRAPH", k + 1)) {
					c.i = k + 2;
					parse_create(c);
					r.handled = true;
					return r;
				}
			}
			if (c.is_word("DROP") && c.is_word("PROPERTY", 1) && c.is_word("GRAPH", 2)) {
				c.i = 3;
				bool if_exists = false;
				if (c.is_word("IF") && c.is_word("EXISTS", 1)) {
					c.i += 2;
					if_exists = true;
				}
				const std::string name = lower(c.ident("a property graph name"));
				if (!graphs_.erase(name) && !if_exists) throw ParseError("property graph " + name + " does not exist");
				r.handled = true;
				return r;
			}
			bool any = false;
			for (size_t k = 0; k + 1 < toks.size(); k++)
				if (toks[k].kind == T_WORD && toks[k].up == "GRAPH_TABLE" && toks[k + 1].kind == T_PUNCT &&
				    toks[k + 1].text[0] == '(')
					any = true;
			if (!any) return r;
			r.handled = true;
			rewrite_query(sql, toks, r);
		} catch (const ParseError &e) {
			r.handled = true;
			r.statements.clear();
			r.error = "pgq: " + e.msg;
		}
		return r;
	}

private:
	std::map<std::string, PropertyGraph> graphs_;
	int next_csr_;

	// [NO PROPERTIES | PROPERTIES ( ... ) | PROPERTIES ARE ALL COLUMNS [EXCEPT ( ... )]] — accepted and ignored: every
	// column of the table stays reachable through the element's variable
	static void skip_properties(detail::Cur &c) {
		if (c.is_word("NO") && c.is_word("PROPERTIES", 1)) {
			c.i += 2;
			return;
		}
		if (!c.accept_word("PROPERTIES")) return;
		if (c.accept_word("ARE")) {
			c.expect_word("ALL");
			c.expect_word("COLUMNS");
			if (c.accept_word("EXCEPT")) c.i = c.match_paren(c.i);
			return;
		}
		if (!c.is_punct('(')) throw detail::ParseError("expected ( after PROPERTIES near " + c.near());
		c.i = c.match_paren(c.i);
	}

	static std::string paren_ident(detail::Cur &c, const char *what) {
		c.expect_punct('(');
		const std::string id = c.ident(what);
		if (c.is_punct(',')) throw detail::ParseError(std::string("composite ") + what + " is not supported");
		c.expect_punct(')');
		return id;
	}

	void parse_create(detail::Cur &c) {
		using namespace detail;
		PropertyGraph g;
		g.name = lower(c.ident("a property graph name"));
		c.expect_word("VERTEX");
		c.expect_word("TABLES");
		c.expect_punct('(');
		for (;;) {
			VertexTable v;
			v.table = c.ident("a vertex table name");
			if (c.accept_word("AS")) v.alias = c.ident("an alias");
			for (;;) {
				if (c.accept_word("KEY")) v.key = paren_ident(c, "key column");
				else if (c.accept_word("LABEL")) v.label = c.ident("a label");
				else if (c.is_word("PROPERTIES") || (c.is_word("NO") && c.is_word("PROPERTIES", 1))) skip_properties(c);
				else break;
			}
			if (v.label.empty()) v.label = v.alias.empty() ? v.table : v.alias;
			g.vertices.push_back(v);
			if (c.accept_punct(',')) continue;
			c.expect_punct(')');
			break;
		}
		if (c.accept_word("EDGE")) {
			c.expect_word("TABLES");
			c.expect_punct('(');
			for (;;) {
				EdgeTable e;
				e.table = c.ident("an edge table name");
				if (c.accept_word("AS")) e.alias = c.ident("an alias");
				for (;;) {
					if (c.accept_word("KEY")) (void)paren_ident(c, "key column");
					else if (c.accept_word("SOURCE")) {
						c.expect_word("KEY");
						e.src_key = paren_ident(c, "source key column");
						c.expect_word("REFERENCES");
						e.src_table = c.ident("a vertex table name");
						e.src_ref = paren_ident(c, "referenced column");
					} else if (c.accept_word("DESTINATION")) {
						c.expect_word("KEY");
						e.dst_key = paren_ident(c, "destination key column");
						c.expect_word("REFERENCES");
						e.dst_table = c.ident("a vertex table name");
						e.dst_ref = paren_ident(c, "referenced column");
					} else if (c.accept_word("LABEL")) e.label = c.ident("a label");
					else if (c.is_word("PROPERTIES") || (c.is_word("NO") && c.is_word("PROPERTIES", 1))) skip_properties(c);
					else break;
				}
				if (e.src_key.empty() || e.dst_key.empty())
					throw ParseError("edge table " + e.table + " needs SOURCE KEY (...) REFERENCES ... and DESTINATION KEY (...) REFERENCES ...");
				if (e.label.empty()) e.label = e.alias.empty() ? e.table : e.alias;
				const VertexTable *s = find_vertex(g, e.src_table), *d = find_vertex(g, e.dst_table);
				if (!s) throw ParseError("edge table " + e.table + " references " + e.src_table + ", which is not a vertex table of " + g.name);
				if (!d) throw ParseError("edge table " + e.table + " references " + e.dst_table + ", which is not a vertex table of " + g.name);
				e.src_table = s->table;
				e.dst_table = d->table;
				g.edges.push_back(e);
				if (c.accept_punct(',')) continue;
				c.expect_punct(')');
				break;
			}
		}
		c.accept_punct(';');
		if (!c.at_end()) throw ParseError("unexpected text after CREATE PROPERTY GRAPH near " + c.near());
		graphs_[g.name] = g;
	}

	struct Piece { // one GRAPH_TABLE ( ... ) occurrence
		size_t tok_b, tok_e; // token span to replace
		std::string subquery;
		bool needs_alias;
	};

	void rewrite_query(const std::string &sql, const std::vector<detail::Tok> &toks, Result &r) {
		using namespace detail;
		std::vector<Piece> pieces;
		std::vector<std::string> ctes;
		std::vector<int> csr_ids;
		Cur all(sql, toks, 0, toks.size());
		for (size_t k = 0; k + 1 < toks.size(); k++) {
			if (!(toks[k].kind == T_WORD && toks[k].up == "GRAPH_TABLE" && toks[k + 1].kind == T_PUNCT &&
			      toks[k + 1].text[0] == '('))
				continue;
			const size_t close = all.match_paren(k + 1); // one past ')'
			Piece p;
			p.tok_b = k;
			p.tok_e = close;
			Cur c(sql, toks, k + 2, close - 1);
			p.subquery = graph_table(c, ctes, csr_ids);
			const Tok &nx = toks[close];
			static const char *const kw[] = {"WHERE", "JOIN",  "ON",    "ORDER",  "GROUP", "LIMIT", "UNION", "HAVING",
			                                 "INNER", "LEFT",  "RIGHT", "FULL",   "CROSS", "NATURAL", "EXCEPT",
			                                 "INTERSECT", "WINDOW", "USING", "OFFSET", "SAMPLE", NULL};
			bool has_alias = nx.kind == T_WORD;
			for (int q = 0; has_alias && kw[q]; q++)
				if (nx.up == kw[q]) has_alias = false;
			p.needs_alias = !has_alias;
			pieces.push_back(p);
			k = close - 1;
		}
		// splice
		std::string out;
		size_t pos = 0;
		// where the CTEs go: the first WITH or SELECT outside parentheses
		size_t ins_tok = toks.size();
		bool ins_with = false;
		int depth = 0;
		for (size_t k = 0; k < toks.size(); k++) {
			if (toks[k].kind == T_PUNCT && toks[k].text[0] == '(') depth++;
			if (toks[k].kind == T_PUNCT && toks[k].text[0] == ')') depth--;
			if (depth == 0 && toks[k].kind == T_WORD && (toks[k].up == "WITH" || toks[k].up == "SELECT")) {
				ins_tok = k;
				ins_with = toks[k].up == "WITH";
				break;
			}
		}
		if (!ctes.empty() && ins_tok == toks.size())
			throw ParseError("GRAPH_TABLE with a path pattern needs a SELECT (or WITH) at the top level of the statement");
		std::string cte_list;
		for (size_t q = 0; q < ctes.size(); q++) cte_list += (q ? ", " : "") + ctes[q];
		size_t next_piece = 0;
		bool inserted = ctes.empty();
		for (size_t k = 0; k < toks.size() && toks[k].kind != T_END; k++) {
			if (!inserted && k == ins_tok) {
				out += sql.substr(pos, toks[k].b - pos);
				pos = toks[k].b;
				if (ins_with) {
					size_t after = k + 1;
					if (toks[after].kind == T_WORD && toks[after].up == "RECURSIVE") after++;
					out += sql.substr(pos, toks[after - 1].e - pos) + " " + cte_list + ", ";
					pos = toks[after - 1].e;
				} else {
					out += "WITH " + cte_list + " ";
				}
				inserted = true;
			}
			if (next_piece < pieces.size() && k == pieces[next_piece].tok_b) {
				const Piece &p = pieces[next_piece];
				out += sql.substr(pos, toks[k].b - pos);
				out += "(" + p.subquery + ")";
				if (p.needs_alias) out += " __pgq_gt" + num((long)next_piece);
				pos = toks[p.tok_e - 1].e;
				k = p.tok_e - 1;
				next_piece++;
			}
		}
		out += sql.substr(pos);
		r.statements.push_back(out);
		r.main_statement = 0;
		for (size_t q = 0; q < csr_ids.size(); q++) r.statements.push_back("SELECT delete_csr(" + num(csr_ids[q]) + ");");
	}

	std::string graph_table(detail::Cur &c, std::vector<std::string> &ctes, std::vector<int> &csr_ids);
};

} // namespace pgq_rewrite

#include "pgq_rewrite_match.hpp"
