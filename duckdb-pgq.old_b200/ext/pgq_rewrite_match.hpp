// GRAPH_TABLE ( graph MATCH pattern [WHERE expr] COLUMNS ( list ) ) -> subquery text (part of pgq_rewrite.hpp).
#pragma once

namespace pgq_rewrite {

inline std::string Rewriter::graph_table(detail::Cur &c, std::vector<std::string> &ctes, std::vector<int> &csr_ids) {
	using namespace detail;
	const std::string gname = lower(c.ident("a property graph name"));
	std::map<std::string, PropertyGraph>::const_iterator git = graphs_.find(gname);
	if (git == graphs_.end()) throw ParseError("property graph " + gname + " does not exist (CREATE PROPERTY GRAPH first)");
	const PropertyGraph &g = git->second;
	c.expect_word("MATCH");

	// ---- pattern ----
	std::string path_var;
	if (c.peek().kind == T_WORD && c.is_punct('=', 1)) {
		path_var = c.ident("a path variable");
		c.expect_punct('=');
	}
	bool any_shortest = false;
	if (c.is_word("ANY") && c.is_word("SHORTEST", 1)) {
		c.i += 2;
		any_shortest = true;
	} else if (c.is_word("ALL") || c.is_word("SHORTEST") || c.is_word("ANY"))
		throw ParseError("only ANY SHORTEST paths are supported, near " + c.near());
	std::vector<PVertex> vs;
	std::vector<PEdge> es;
	std::vector<std::string> vlabels, elabels; // as written ("" = none)
	int anon = 0;
	for (;;) {
		// vertex: ( var [: Label] [WHERE expr] )
		if (!c.is_punct('(')) throw ParseError("expected a vertex pattern '(' near " + c.near());
		const size_t vclose = c.match_paren(c.i);
		Cur vc(c.src, c.t, c.i + 1, vclose - 1);
		PVertex v;
		v.tbl = NULL;
		std::string label;
		if (vc.peek().kind == T_WORD && !vc.is_word("WHERE")) v.var = vc.ident("a vertex variable");
		if (vc.accept_punct(':')) label = vc.ident("a vertex label");
		if (vc.accept_word("WHERE")) {
			v.where = span_text(c.src, c.t, vc.i, vc.end);
			vc.i = vc.end;
		}
		if (!vc.at_end() && vc.i < vc.end) throw ParseError("unexpected text in a vertex pattern near " + vc.near());
		if (v.var.empty()) v.var = "__pgq_v" + num(anon++);
		vs.push_back(v);
		vlabels.push_back(label);
		c.i = vclose;
		// edge?
		PEdge e;
		if (c.is_punct('-') && c.is_punct('[', 1)) {
			c.i += 2;
		} else if (c.is_punct('<') && c.is_punct('-', 1) && c.is_punct('[', 2)) {
			c.i += 3;
			e.reversed = true;
		} else
			break;
		std::string elabel;
		if (c.peek().kind == T_WORD) e.var = c.ident("an edge variable");
		if (c.accept_punct(':')) elabel = c.ident("an edge label");
		if (c.is_word("WHERE")) throw ParseError("WHERE inside an edge pattern is not supported: put the predicate after the pattern");
		c.expect_punct(']');
		c.expect_punct('-');
		if (!e.reversed) c.expect_punct('>');
		else if (c.is_punct('>')) throw ParseError("undirected / bidirectional edge patterns are not supported");
		if (c.accept_punct('*')) {
			e.quantified = true;
		} else if (c.accept_punct('+')) {
			e.quantified = true;
			e.lower = 1;
		} else if (c.accept_punct('{')) {
			e.quantified = true;
			if (c.peek().kind == T_NUMBER) e.lower = atol(c.t[c.i++].text.c_str());
			if (c.accept_punct(',')) {
				if (c.peek().kind == T_NUMBER) e.upper = atol(c.t[c.i++].text.c_str());
			} else
				e.upper = e.lower;
			c.expect_punct('}');
			if (e.upper >= 0 && e.upper < e.lower) throw ParseError("path quantifier with upper < lower");
		}
		if (e.var.empty()) e.var = "__pgq_e" + num(anon++);
		es.push_back(e);
		elabels.push_back(elabel);
	}
	if (es.empty()) throw ParseError("a MATCH pattern needs at least one edge, near " + c.near());

	// ---- resolve labels ----
	for (size_t k = 0; k < es.size(); k++) {
		if (!elabels[k].empty()) {
			es[k].tbl = find_edge(g, elabels[k]);
			if (!es[k].tbl) throw ParseError("no edge table with label " + elabels[k] + " in property graph " + gname);
		} else if (g.edges.size() == 1) {
			es[k].tbl = &g.edges[0];
		} else
			throw ParseError("edge pattern " + es[k].var + " needs a label: property graph " + gname + " has several edge tables");
	}
	for (size_t k = 0; k < vs.size(); k++) {
		if (!vlabels[k].empty()) {
			vs[k].tbl = find_vertex(g, vlabels[k]);
			if (!vs[k].tbl) throw ParseError("no vertex table with label " + vlabels[k] + " in property graph " + gname);
			continue;
		}
		std::string want; // from the adjacent edges
		if (k < es.size()) want = es[k].reversed ? es[k].tbl->dst_table : es[k].tbl->src_table;
		else want = es[k - 1].reversed ? es[k - 1].tbl->src_table : es[k - 1].tbl->dst_table;
		vs[k].tbl = find_vertex(g, want);
		if (!vs[k].tbl) throw ParseError("cannot infer the vertex table of " + vs[k].var);
	}
	for (size_t k = 0; k < es.size(); k++) {
		const PVertex &from = es[k].reversed ? vs[k + 1] : vs[k], &to = es[k].reversed ? vs[k] : vs[k + 1];
		if (!same_ident(from.tbl->table, es[k].tbl->src_table) || !same_ident(to.tbl->table, es[k].tbl->dst_table))
			throw ParseError("edge " + es[k].var + " (" + es[k].tbl->table + ") runs from " + es[k].tbl->src_table + " to " +
			                 es[k].tbl->dst_table + ", not from " + from.tbl->table + " to " + to.tbl->table);
		if (es[k].quantified && !same_ident(es[k].tbl->src_table, es[k].tbl->dst_table))
			throw ParseError("a path over " + es[k].tbl->table + " needs source and destination in the same vertex table");
	}
	size_t path_edge = es.size(); // the (single) quantified edge the path variable names
	for (size_t k = 0; k < es.size(); k++)
		if (es[k].quantified) {
			if (path_edge != es.size() && !path_var.empty())
				throw ParseError("a path variable over more than one quantified edge is not supported");
			path_edge = k;
		}
	if (!path_var.empty() && path_edge == es.size())
		throw ParseError("path variable " + path_var + " needs a quantified edge (->*, ->+ or ->{l,u})");
	if (any_shortest && path_edge == es.size()) throw ParseError("ANY SHORTEST needs a quantified edge");

	// ---- optional WHERE, then COLUMNS ( ... ) ----
	size_t where_b = 0, where_e = 0;
	if (c.accept_word("WHERE")) {
		where_b = c.i;
		int depth = 0;
		while (!c.at_end() && !(depth == 0 && c.is_word("COLUMNS") && c.is_punct('(', 1))) {
			if (c.is_punct('(')) depth++;
			if (c.is_punct(')')) depth--;
			c.i++;
		}
		where_e = c.i;
	}
	c.expect_word("COLUMNS");
	if (!c.is_punct('(')) throw ParseError("expected ( after COLUMNS near " + c.near());
	const size_t cols_close = c.match_paren(c.i);
	const size_t cols_b = c.i + 1, cols_e = cols_close - 1;
	c.i = cols_close;
	if (!c.at_end() && c.i < c.end) throw ParseError("unexpected text after COLUMNS ( ... ) near " + c.near());

	// ---- path function calls ----
	// call expressions of quantified edge k (source side first)
	struct Calls {
		// |V|: a column of the one-row aggregate, not a scalar subquery — with a subquery inside a WHERE predicate this
		// engine plans a blockwise nested-loop join and calls the function one row at a time (71 us per row, measured in
		// round 1: profiles/r01_sql_path.md)
		static std::string cnt(const PEdge &e) {
			return "__pgq_x" + num(e.csr_id) + ".n";
		}
		static std::string sd(const PEdge &e, const PVertex &a, const PVertex &b) {
			const PVertex &s = e.reversed ? b : a, &d = e.reversed ? a : b;
			return s.var + ".rowid + __pgq_x" + num(e.csr_id) + ".z, " + d.var + ".rowid";
		}
		static std::string bounds(const PEdge &e) {
			if (e.lower == 0 && e.upper < 0) return "";
			return ", CAST(" + num(e.lower) + " AS BIGINT), " +
			       (e.upper < 0 ? cnt(e) : "CAST(" + num(e.upper) + " AS BIGINT)");
		}
		static std::string length(const PEdge &e, const PVertex &a, const PVertex &b) {
			return "iterativelength(" + num(e.csr_id) + ", " + cnt(e) + ", " + sd(e, a, b) + bounds(e) + ")";
		}
	};
	for (size_t k = 0; k < es.size(); k++)
		if (es[k].quantified) {
			es[k].csr_id = next_csr_++;
			if (next_csr_ > (1 << 30)) next_csr_ = 1000;
		}
	bool used_path_fn = false;
	// text of tokens [b, e) with path-function calls replaced
	struct Subst {
		static std::string run(const std::string &src, const std::vector<Tok> &t, size_t b, size_t e, const std::string &pv,
		                       std::vector<PEdge> &es, const std::vector<PVertex> &vs, size_t pe, bool &used) {
			std::string out;
			size_t pos = b < e ? t[b].b : 0;
			for (size_t k = b; k < e; k++) {
				if (t[k].kind != T_WORD || k + 2 >= e || !(t[k + 1].kind == T_PUNCT && t[k + 1].text[0] == '(')) continue;
				const std::string &f = t[k].up;
				const bool is_fn = f == "PATH_LENGTH" || f == "VERTICES" || f == "ELEMENT_ID" || f == "PATH_WITH_EDGES" ||
				                   f == "CHEAPEST_PATH_LENGTH" || f == "CHEAPEST_PATH_LENGTH_DOUBLE";
				if (!is_fn) continue;
				if (!(t[k + 2].kind == T_WORD && !pv.empty() && same_ident(t[k + 2].text, pv))) continue;
				// find the closing parenthesis
				int depth = 0;
				size_t close = k + 1;
				for (; close < e; close++) {
					if (t[close].kind == T_PUNCT && t[close].text[0] == '(') depth++;
					if (t[close].kind == T_PUNCT && t[close].text[0] == ')' && --depth == 0) break;
				}
				if (close >= e) throw ParseError("unbalanced parentheses in " + t[k].text + "(...)");
				PEdge &pe_ = es[pe];
				const PVertex &a = vs[pe], &bq = vs[pe + 1];
				std::string call;
				if (f == "PATH_LENGTH") call = Calls::length(pe_, a, bq);
				else if (f == "VERTICES" || f == "ELEMENT_ID")
					call = "shortestpath(" + num(pe_.csr_id) + ", " + Calls::cnt(pe_) + ", " + Calls::sd(pe_, a, bq) + ")";
				else if (f == "PATH_WITH_EDGES")
					call = "shortestpath_with_edges(" + num(pe_.csr_id) + ", " + Calls::cnt(pe_) + ", " + Calls::sd(pe_, a, bq) + ")";
				else { // cheapest_path_length[_double](p, edgevar.col)
					if (!(k + 6 <= close && t[k + 3].kind == T_PUNCT && t[k + 3].text[0] == ',' && t[k + 4].kind == T_WORD &&
					      t[k + 5].kind == T_PUNCT && t[k + 5].text[0] == '.' && t[k + 6].kind == T_WORD && close == k + 7))
						throw ParseError(t[k].text + " takes (path variable, edge_variable.weight_column)");
					if (!same_ident(t[k + 4].text, pe_.var))
						throw ParseError("the weight column of " + t[k].text + " must belong to the path's edge variable " + pe_.var);
					const bool dbl = f == "CHEAPEST_PATH_LENGTH_DOUBLE";
					if (!pe_.weight_col.empty() && (!same_ident(pe_.weight_col, t[k + 6].text) || pe_.weight_double != dbl))
						throw ParseError("one weight column per path");
					pe_.weight_col = t[k + 6].text;
					pe_.weight_double = dbl;
					call = std::string(dbl ? "cheapest_path_length_double(" : "cheapest_path_length(") + num(pe_.csr_id) + ", " +
					       Calls::cnt(pe_) + ", " + Calls::sd(pe_, a, bq) + ")";
				}
				out += src.substr(pos, t[k].b - pos) + call;
				pos = t[close].e;
				k = close;
				used = true;
			}
			if (b < e) out += src.substr(pos, t[e - 1].e - pos);
			return out;
		}
	};
	std::string cols = Subst::run(c.src, c.t, cols_b, cols_e, path_var, es, vs, path_edge == es.size() ? 0 : path_edge, used_path_fn);
	std::string where = where_b < where_e
	                        ? Subst::run(c.src, c.t, where_b, where_e, path_var, es, vs, path_edge == es.size() ? 0 : path_edge, used_path_fn)
	                        : std::string();
	{ // COLUMNS (*)
		Cur cc(c.src, c.t, cols_b, cols_e);
		if (cc.is_punct('*') && cols_e == cols_b + 1) {
			cols.clear();
			for (size_t k = 0; k < vs.size(); k++) cols += (cols.empty() ? "" : ", ") + vs[k].var + ".*";
			for (size_t k = 0; k < es.size(); k++)
				if (!es[k].quantified) cols += ", " + es[k].var + ".*";
		}
	}

	// ---- CSR-building CTEs (one per quantified edge) ----
	for (size_t k = 0; k < es.size(); k++) {
		const PEdge &e = es[k];
		if (!e.quantified) continue;
		const EdgeTable &et = *e.tbl;
		const VertexTable *vt = find_vertex(g, et.src_table);
		const std::string id = num(e.csr_id), VT = vt->table, cnt = "(SELECT count(*) FROM " + VT + ")";
		std::string q = "__pgq_csr" + id + " AS (SELECT create_csr_edge(" + id + ", " + cnt +
		                ", CAST((SELECT sum(create_csr_vertex(" + id + ", " + cnt + ", sub.dense_id, sub.cnt)) FROM (SELECT v.rowid AS dense_id, "
		                "count(e.sk) AS cnt FROM " + VT + " v LEFT JOIN (SELECT e." + et.src_key + " AS sk FROM " + et.table + " e JOIN " + VT +
		                " d ON d." + et.dst_ref + " = e." + et.dst_key + ") e ON e.sk = v." + et.src_ref +
		                " GROUP BY v.rowid) sub) AS BIGINT), s.rowid, d.rowid, e.rowid" +
		                (e.weight_col.empty() ? std::string()
		                                      : ", CAST(e." + e.weight_col + (e.weight_double ? " AS DOUBLE)" : " AS BIGINT)")) +
		                ") AS temp FROM " + et.table + " e JOIN " + VT + " s ON s." + et.src_ref + " = e." + et.src_key + " JOIN " + VT +
		                " d ON d." + et.dst_ref + " = e." + et.dst_key + ")";
		ctes.push_back(q);
		csr_ids.push_back(e.csr_id);
	}

	// ---- the subquery ----
	std::string from, cond;
	for (size_t k = 0; k < es.size(); k++)
		if (es[k].quantified) {
			const std::string id = num(es[k].csr_id);
			from += (from.empty() ? "" : ", ") + std::string("(SELECT count(__pgq_csr") + id + ".temp) * 0 AS z, (SELECT count(*) FROM " +
			        es[k].tbl->src_table + ") AS n FROM __pgq_csr" + id + ") __pgq_x" + id;
		}
	for (size_t k = 0; k < vs.size(); k++) from += (from.empty() ? "" : ", ") + vs[k].tbl->table + " " + vs[k].var;
	for (size_t k = 0; k < es.size(); k++)
		if (!es[k].quantified) from += ", " + es[k].tbl->table + " " + es[k].var;
	for (size_t k = 0; k < es.size(); k++) {
		const PEdge &e = es[k];
		const PVertex &s = e.reversed ? vs[k + 1] : vs[k], &d = e.reversed ? vs[k] : vs[k + 1];
		std::string one;
		if (!e.quantified)
			one = e.var + "." + e.tbl->src_key + " = " + s.var + "." + e.tbl->src_ref + " AND " + e.var + "." + e.tbl->dst_key + " = " +
			      d.var + "." + e.tbl->dst_ref;
		else if (k == path_edge && (used_path_fn || any_shortest))
			one = Calls::length(e, vs[k], vs[k + 1]) + " IS NOT NULL";
		else if (e.lower == 0 && e.upper < 0)
			one = "reachability(" + num(e.csr_id) + ", " + Calls::cnt(e) + ", " + Calls::sd(e, vs[k], vs[k + 1]) + ")";
		else
			one = Calls::length(e, vs[k], vs[k + 1]) + " IS NOT NULL";
		cond += (cond.empty() ? "" : " AND ") + one;
	}
	for (size_t k = 0; k < vs.size(); k++)
		if (!vs[k].where.empty()) cond += " AND (" + vs[k].where + ")";
	if (!where.empty()) cond += " AND (" + where + ")";
	return "SELECT " + cols + " FROM " + from + " WHERE " + cond;
}

} // namespace pgq_rewrite
