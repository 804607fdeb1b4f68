.timer on
SET threads=16;
CREATE TABLE t AS SELECT i AS id, [((hash(i * 128 + x) % 100000) / 100000.0)::FLOAT for x in range(128)]::FLOAT[128] AS vec FROM range(1000000) r(i);
CREATE TABLE q AS SELECT i AS qid, [((hash((i + 77777) * 128 + x) % 100000) / 100000.0)::FLOAT for x in range(128)]::FLOAT[128] AS qv FROM range(256) r(i);
SET hnsw_exact_search = true;
.print ##join
SELECT count(*), sum(id) FROM q, LATERAL (SELECT id FROM t ORDER BY array_distance(t.vec, q.qv) LIMIT 10);
.print ##join_again
SELECT count(*), sum(id) FROM q, LATERAL (SELECT id FROM t ORDER BY array_distance(t.vec, q.qv) LIMIT 10);
.print ##count_scan
SELECT count(*), sum(vec[1]) FROM t;
