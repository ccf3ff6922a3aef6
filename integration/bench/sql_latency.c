/* sql_latency.c — SQL-level latency of the rewritten kNN query through the vendored DuckDB with the patched extension over
 * liblance_b200 (integration/build_duckdb_harness.sh stages libduckdb.so).  DuckDB C API only.
 *   gcc -O2 integration/bench/sql_latency.c -I /root/reference/duckdb/src/include -L integration/_build -lduckdb \
 *       -Wl,-rpath,'$ORIGIN' -o integration/_build/sql_latency
 *   integration/_build/sql_latency [rows] [dim] [queries]
 * Prints the mean wall-clock per `SELECT id FROM t ORDER BY array_distance(v, q) LIMIT 10` (parse + plan + rewrite +
 * LANCE_INDEX_SCAN + fetch) and per `lance_search(...)` table-function call, with and without the LANCE index. */
#include <duckdb.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

static double now_ms(void) {
	struct timespec ts;
	clock_gettime(CLOCK_MONOTONIC, &ts);
	return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6;
}

static void run(duckdb_connection con, const char *sql) {
	duckdb_result r;
	if (duckdb_query(con, sql, &r) == DuckDBError) {
		fprintf(stderr, "query failed: %s\n  %s\n", duckdb_result_error(&r), sql);
		exit(1);
	}
	duckdb_destroy_result(&r);
}

static double time_query(duckdb_connection con, const char *sql, int reps, long long *first_id) {
	duckdb_result r;
	for (int i = 0; i < 3; i++) {
		if (duckdb_query(con, sql, &r) == DuckDBError) {
			fprintf(stderr, "query failed: %s\n", duckdb_result_error(&r));
			exit(1);
		}
		if (i == 0 && first_id) *first_id = duckdb_row_count(&r) ? duckdb_value_int64(&r, 0, 0) : -1;
		duckdb_destroy_result(&r);
	}
	const double t0 = now_ms();
	for (int i = 0; i < reps; i++) {
		duckdb_query(con, sql, &r);
		duckdb_destroy_result(&r);
	}
	return (now_ms() - t0) / reps;
}

int main(int argc, char **argv) {
	const long long rows = argc > 1 ? atoll(argv[1]) : 100000;
	const int dim = argc > 2 ? atoi(argv[2]) : 128;
	const int reps = argc > 3 ? atoi(argv[3]) : 200;
	duckdb_database db;
	duckdb_connection con;
	if (duckdb_open(NULL, &db) == DuckDBError || duckdb_connect(db, &con) == DuckDBError) return 1;
	char *sql = malloc(64 * 1024);
	run(con, "SELECT setseed(0.42)");
	snprintf(sql, 65536,
	         "CREATE TABLE t AS SELECT i AS id, list_transform(range(%d), x -> (random() * 2 - 1)::FLOAT)::FLOAT[%d] AS v "
	         "FROM range(%lld) r(i)", dim, dim, rows);
	double t0 = now_ms();
	run(con, sql);
	printf("table %lld x %d built in %.0f ms\n", rows, dim, now_ms() - t0);
	/* the query vector as a literal */
	char *q = malloc(32 * 1024);
	int o = sprintf(q, "[");
	for (int i = 0; i < dim; i++) o += sprintf(q + o, "%s%.4f", i ? "," : "", (double)((i * 37 % 100) - 50) / 50.0);
	sprintf(q + o, "]::FLOAT[%d]", dim);
	long long id_scan = -1, id_index = -1, id_fn = -1;
	snprintf(sql, 65536, "SELECT id FROM t ORDER BY array_distance(v, %s) LIMIT 10", q);
	const double ms_scan = time_query(con, sql, reps > 20 ? 20 : reps, &id_scan);
	t0 = now_ms();
	run(con, "CREATE INDEX t_idx ON t USING LANCE (v)");
	printf("CREATE INDEX ... USING LANCE: %.0f ms\n", now_ms() - t0);
	const double ms_index = time_query(con, sql, reps, &id_index);
	char *sql2 = malloc(64 * 1024);
	snprintf(sql2, 65536, "SELECT * FROM lance_search('t', 't_idx', %s, 10)", q);
	/* lance_search returns row ids; only its latency is compared */
	const double ms_fn = time_query(con, sql2, reps, &id_fn);
	printf("{\"rows\": %lld, \"dim\": %d, \"order_by_limit_no_index_ms\": %.3f, \"order_by_limit_lance_index_ms\": %.3f, "
	       "\"lance_search_ms\": %.3f, \"same_first_id\": %s}\n",
	       rows, dim, ms_scan, ms_index, ms_fn, id_scan == id_index ? "true" : "false");
	duckdb_disconnect(&con);
	duckdb_close(&db);
	return 0;
}
