// pgq_rewrite_cli — the SQL/PGQ rewriter (pgq_rewrite.hpp) on its own: statements on stdin (a ';' at the end of a line
// ends one), the rewritten statements on stdout.  No DuckDB, no GPU: what the CPU tests and a curious reader run.
//   "-- passthrough N" + the statement | "-- rewritten N: k statements, main m" + the statements | "-- error N: ..."
#include "pgq_rewrite.hpp"

#include <iostream>

int main() {
	pgq_rewrite::Rewriter rw;
	std::string line, stmt;
	int n = 0;
	while (std::getline(std::cin, line)) {
		stmt += line;
		stmt += "\n";
		size_t end = line.find_last_not_of(" \t\r");
		if (end == std::string::npos || line[end] != ';') continue;
		n++;
		pgq_rewrite::Result r = rw.rewrite(stmt);
		if (!r.handled) std::cout << "-- passthrough " << n << "\n" << stmt;
		else if (!r.error.empty()) std::cout << "-- error " << n << ": " << r.error << "\n";
		else {
			std::cout << "-- rewritten " << n << ": " << r.statements.size() << " statements, main " << r.main_statement << "\n";
			for (size_t k = 0; k < r.statements.size(); k++) {
				std::cout << r.statements[k];
				if (r.statements[k].empty() || r.statements[k][r.statements[k].size() - 1] != '\n') std::cout << "\n";
			}
		}
		stmt.clear();
	}
	std::cout << "-- done " << n << "\n";
	return 0;
}
