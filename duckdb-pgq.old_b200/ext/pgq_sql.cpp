// pgq_sql — minimal SQL driver over the reference DuckDB's C++ API (Connection::Query,
// /root/reference/src/include/duckdb/main/connection.hpp).  Reads statements from stdin (separated by
// ';' at end of line), prints each result row as comma-separated values after a "-- result N" line,
// and errors as "-- error N: <message>".  It links the same libduckdb.so as the extension, so the
// plugin and the host share one copy of DuckDB (SURVEY.md §8(b) caveat about the static shell).
#include "duckdb.hpp"
#include "pgq_rewrite.hpp"

#include <chrono>
#include <cstdlib>
#include <iostream>
#include <sstream>
#include <string>
#include <vector>

using namespace duckdb;

int main(int argc, char **argv) {
	DuckDB db(argc > 1 ? argv[1] : nullptr);
	Connection con(db);
	pgq_rewrite::Rewriter rewriter;
	std::string line, stmt;
	int n = 0, failed = 0;
	const bool timing = std::getenv("PGQ_SQL_TIME") != nullptr;
	while (std::getline(std::cin, line)) {
		stmt += line;
		stmt += "\n";
		size_t end = line.find_last_not_of(" \t\r");
		if (end == std::string::npos || line[end] != ';') continue;
		// SQL/PGQ surface syntax (CREATE PROPERTY GRAPH, GRAPH_TABLE ... MATCH) becomes plain SQL over the path's
		// scalar functions first (pgq_rewrite.hpp; PGQ_SQL_SHOW_REWRITE=1 prints what is run instead)
		pgq_rewrite::Result rw = rewriter.rewrite(stmt);
		std::vector<std::string> after;
		if (rw.handled) {
			n++;
			stmt.clear();
			if (!rw.error.empty()) {
				std::cout << "-- error " << n << ": " << rw.error << "\n";
				failed++;
				continue;
			}
			if (rw.statements.empty()) {
				std::cout << "-- result " << n << "\n";
				continue;
			}
			if (std::getenv("PGQ_SQL_SHOW_REWRITE"))
				for (auto &q : rw.statements) std::cout << "-- rewrite " << n << ": " << q << "\n";
			stmt = rw.statements[rw.main_statement];
			for (size_t k = rw.main_statement + 1; k < rw.statements.size(); k++) after.push_back(rw.statements[k]);
			n--;
		}
		auto t0 = std::chrono::steady_clock::now();
		auto res = con.Query(stmt);
		auto t1 = std::chrono::steady_clock::now();
		for (auto &q : after) con.Query(q); // delete_csr of the CSRs the statement built
		stmt.clear();
		n++;
		if (timing) // PGQ_SQL_TIME=1: wall time of Connection::Query (parse + plan + execute + materialise)
			std::cout << "-- time " << n << ": " << std::chrono::duration<double, std::milli>(t1 - t0).count() << " ms\n";
		if (!res->success) {
			std::cout << "-- error " << n << ": " << res->error << "\n";
			failed++;
			continue;
		}
		std::cout << "-- result " << n << "\n";
		for (idx_t r = 0; r < res->collection.Count(); r++) {
			for (idx_t c = 0; c < res->ColumnCount(); c++) {
				if (c) std::cout << ",";
				Value v = res->GetValue(c, r);
				std::cout << (v.is_null ? std::string("NULL") : v.ToString());
			}
			std::cout << "\n";
		}
	}
	std::cout << "-- done " << n << " statements, " << failed << " errors\n";
	return 0;
}
