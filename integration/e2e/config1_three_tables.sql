-- BASELINE config 1 as SURVEY §8(d) prescribes it: the README tables (README.md:70-83) plus a third table, because the
-- optimizer leaves a 2-table join untouched (≤ 1 join edge, src/optimizer/rpt_optimizer.cpp:1592-1595) and only the
-- 3-table variant plans CREATE_BF / USE_BF.  Also the survey's a ⋈ b ⋈ c shape with single-column keys.
CREATE TABLE t1 AS SELECT range AS id, range * 3 AS v FROM range(100000);
CREATE TABLE t2 AS SELECT range AS id, range % 7 AS w FROM range(50000);
CREATE TABLE t3 AS SELECT range * 2 AS id FROM range(20000);
SELECT 'readme_2table', count(*) FROM t1 JOIN t2 USING (id);
SELECT 'readme_3table', count(*), sum(t1.v), sum(t2.w) FROM t1 JOIN t2 USING (id) JOIN t3 USING (id);
CREATE TABLE big AS SELECT range AS k, hash(range) % 16667 AS fk1, hash(range + 7) % 100 AS fk2 FROM range(1000000);
CREATE TABLE mid AS SELECT range AS id, range % 13 AS tag FROM range(16667) WHERE range % 3 = 0;
CREATE TABLE small AS SELECT range AS id FROM range(100) WHERE range % 4 = 1;
SELECT 'toy_1m_16667_100', count(*), sum(big.k), sum(mid.tag) FROM big JOIN mid ON big.fk1 = mid.id JOIN small ON big.fk2 = small.id;
-- string key: VARCHAR filter through the operators
CREATE TABLE names AS SELECT range AS id, 'name-' || (range * 7919 % 100003)::VARCHAR AS nm FROM range(100003) WHERE range % 5 = 0;
CREATE TABLE refs AS SELECT range AS rid, 'name-' || (hash(range) % 100003)::VARCHAR AS nm, hash(range) % 100 AS fk2 FROM range(400000);
SELECT 'varchar_key', count(*), sum(refs.rid) FROM refs JOIN names USING (nm) JOIN small ON refs.fk2 = small.id;
-- empty build side: USE_BF emits nothing (physical_use_bf.cpp:145-155)
CREATE TABLE nothing AS SELECT range AS id FROM range(10) WHERE range > 100;
SELECT 'empty_build', count(*) FROM big JOIN mid ON big.fk1 = mid.id JOIN nothing ON big.fk2 = nothing.id;
-- NULL keys never join; the filter inserts and probes them like values (vector_hash.cpp:17-22)
CREATE TABLE withnull AS SELECT CASE WHEN range % 10 = 0 THEN NULL ELSE range END AS id FROM range(30000);
SELECT 'null_keys', count(*) FROM t1 JOIN withnull USING (id) JOIN t3 USING (id);
