// Sanitizer driver for the host runtime: links csrc/host/*.cc directly and drives the C ABI the way the
// Python layer does — multi-threaded text load with key collection, batch dealing, concurrent packing from
// several threads, archive and pass-file round trips, host table bulk ops, SSD spill / promote, key sort.
// Built three times by benchmarks/sanitize_host.sh: -fsanitize=address,undefined, -fsanitize=thread, plain.
#include <algorithm>
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../../../include/pbx_host.h"

#define OK(call)                                                                          \
  do {                                                                                    \
    if ((call) != 0) {                                                                    \
      fprintf(stderr, "FAILED %s:%d %s: %s\n", __FILE__, __LINE__, #call, pbh_last_error()); \
      return 1;                                                                           \
    }                                                                                     \
  } while (0)
#define CHECK(c)                                                       \
  do {                                                                 \
    if (!(c)) { fprintf(stderr, "FAILED %s:%d %s\n", __FILE__, __LINE__, #c); return 1; } \
  } while (0)

static void* make_dataset(int S) {
  void* ds = nullptr;
  pbh_dataset_create(&ds);
  std::vector<std::string> names = {"label", "dense"};
  std::vector<int> types = {1, 1}, dense = {1, 1}, used = {1, 1}, dims = {1, 3};
  for (int s = 0; s < S; ++s) { names.push_back("s" + std::to_string(s)); types.push_back(0); dense.push_back(0); used.push_back(1); dims.push_back(1); }
  std::vector<const char*> np;
  for (auto& n : names) np.push_back(n.c_str());
  pbh_dataset_set_slots(ds, (int)np.size(), np.data(), types.data(), dense.data(), used.data(), dims.data());
  return ds;
}

int main(int argc, char** argv) {
  const std::string dir = argc > 1 ? argv[1] : "/tmp";
  const int S = 6, FILES = 4, LINES = 3000;
  std::vector<std::string> files;
  unsigned rng = 12345;
  auto rnd = [&] { rng = rng * 1103515245u + 12345u; return (rng >> 8) & 0xffffff; };
  for (int f = 0; f < FILES; ++f) {
    files.push_back(dir + "/san-" + std::to_string(f) + ".txt");
    FILE* fp = fopen(files.back().c_str(), "w");
    CHECK(fp != nullptr);
    for (int i = 0; i < LINES; ++i) {
      fprintf(fp, "1 %d 3 0.5 0.25 %g", (int)(rnd() & 1), (rnd() % 100) / 100.0 + 0.01);
      for (int s = 0; s < S; ++s) {
        int n = 1 + rnd() % 3;
        fprintf(fp, " %d", n);
        for (int k = 0; k < n; ++k) fprintf(fp, " %u", 1 + rnd() % 5000 + s * 5000);
      }
      fputc('\n', fp);
    }
    fclose(fp);
  }
  std::vector<const char*> fptr;
  for (auto& f : files) fptr.push_back(f.c_str());

  void* ds = make_dataset(S);
  OK(pbh_dataset_set_filelist(ds, FILES, fptr.data()));
  OK(pbh_dataset_set_int(ds, "thread_num", 4));
  OK(pbh_dataset_set_int(ds, "merge_thread_num", 3));
  OK(pbh_dataset_set_int(ds, "batch_size", 128));
  OK(pbh_dataset_preload_into_memory(ds));
  OK(pbh_dataset_wait_preload_done(ds));
  CHECK(pbh_dataset_num_records(ds) == FILES * LINES);
  int64_t nuniq = pbh_dataset_unique_keys(ds, nullptr, 0);
  CHECK(nuniq > 1000);
  OK(pbh_dataset_prepare_train(ds, 3, 1, 7, 0));
  // concurrent packing: 3 workers x 2 threads each, interleaved batches
  std::atomic<int> bad{0};
  std::atomic<long long> key_sum{0};
  std::vector<std::thread> packers;
  for (int w = 0; w < 3; ++w)
    for (int t = 0; t < 2; ++t)
      packers.emplace_back([&, w, t] {
        int64_t nb = pbh_dataset_num_batches(ds, w);
        for (int64_t b = t; b < nb; b += 2) {
          int B = 0; int64_t nk = 0;
          if (pbh_dataset_batch_shape(ds, w, b, &B, &nk) != 0) { ++bad; return; }
          std::vector<uint64_t> keys((size_t)nk + 1);
          std::vector<int32_t> off((size_t)B * (S + 1) + S * B + 2);
          std::vector<float> dense((size_t)B * 4);
          if (pbh_dataset_pack_batch(ds, w, b, keys.data(), off.data(), dense.data(), (int)(b & 1)) != 0) { ++bad; return; }
          long long s = 0;
          for (int64_t i = 0; i < nk; ++i) s += (long long)keys[i];
          key_sum += s;
        }
      });
  for (auto& th : packers) th.join();
  CHECK(bad.load() == 0 && key_sum.load() > 0);
  // archive + pass file round trips
  const std::string arc = dir + "/san.archive", arena = dir + "/san.arena";
  OK(pbh_dataset_save_archive(ds, arc.c_str()));
  OK(pbh_dataset_save_arena(ds, arena.c_str()));
  void* ds2 = make_dataset(S);
  OK(pbh_dataset_set_int(ds2, "batch_size", 128));
  OK(pbh_dataset_load_archive(ds2, arc.c_str()));
  CHECK(pbh_dataset_num_records(ds2) == FILES * LINES);
  OK(pbh_dataset_load_arena(ds2, arena.c_str()));
  CHECK(pbh_dataset_num_records(ds2) == FILES * LINES && pbh_dataset_unique_keys(ds2, nullptr, 0) == nuniq);
  OK(pbh_dataset_prepare_train(ds2, 2, 1, 3, 0));
  {
    int B = 0; int64_t nk = 0;
    OK(pbh_dataset_batch_shape(ds2, 1, 0, &B, &nk));
    std::vector<uint64_t> keys((size_t)nk + 1);
    std::vector<int32_t> off((size_t)B * (S + 1) + S * B + 2);
    std::vector<float> dense((size_t)B * 4);
    OK(pbh_dataset_pack_batch(ds2, 1, 0, keys.data(), off.data(), dense.data(), 0));
    OK(pbh_dataset_pack_batch(ds2, 1, 0, keys.data(), off.data(), dense.data(), 1));
  }
  CHECK(pbh_dataset_slots_shuffle(ds2, nullptr, 0, 1) != 0);          // refused on a pass file
  // damaged archives and pass files: truncated at many lengths, bytes flipped — an error or a clean load, never a
  // read outside the mapping
  {
    auto slurp = [](const std::string& p) {
      std::vector<char> v;
      FILE* f = fopen(p.c_str(), "rb");
      if (!f) return v;
      fseek(f, 0, SEEK_END);
      long n = ftell(f);
      fseek(f, 0, SEEK_SET);
      v.resize((size_t)n);
      if (n && fread(v.data(), 1, (size_t)n, f) != (size_t)n) v.clear();
      fclose(f);
      return v;
    };
    auto spit = [](const std::string& p, const char* d, size_t n) {
      FILE* f = fopen(p.c_str(), "wb");
      if (n) fwrite(d, 1, n, f);
      fclose(f);
    };
    const std::vector<char> a0 = slurp(arc), p0 = slurp(arena);
    CHECK(!a0.empty() && !p0.empty());
    const std::string bad = dir + "/san.damaged";
    void* dd = make_dataset(S);
    OK(pbh_dataset_set_int(dd, "batch_size", 64));
    for (int t = 0; t < 60; ++t) {
      const std::vector<char>& src = (t & 1) ? p0 : a0;
      std::vector<char> v = src;
      if (t % 3 == 0) v.resize(rnd() % v.size());                                   // truncated
      else for (int k = 0; k < 8; ++k) v[rnd() % std::min<size_t>(v.size(), 4096)] ^= (char)(1 + rnd() % 255);   // header / early bytes flipped
      spit(bad, v.data(), v.size());
      int rc = (t & 1) ? pbh_dataset_load_arena(dd, bad.c_str()) : pbh_dataset_load_archive(dd, bad.c_str());
      if (rc == 0) {                              // a damaged file may still be self-consistent: it must then be usable
        if (pbh_dataset_prepare_train(dd, 1, 0, 0, 0) == 0 && pbh_dataset_num_batches(dd, 0) > 0) {
          int B = 0; int64_t nk = 0;
          if (pbh_dataset_batch_shape(dd, 0, 0, &B, &nk) == 0 && nk >= 0 && nk < (1 << 26)) {
            std::vector<uint64_t> keys((size_t)nk + 1);
            std::vector<int32_t> off((size_t)B * (S + 1) + S * B + 2);
            std::vector<float> dense((size_t)B * 4);
            pbh_dataset_pack_batch(dd, 0, 0, keys.data(), off.data(), dense.data(), 0);
          }
        }
      }
      pbh_dataset_release_memory(dd);
    }
    OK(pbh_dataset_destroy(dd));
  }
  OK(pbh_dataset_release_memory(ds2));
  OK(pbh_dataset_destroy(ds2));
  std::vector<uint64_t> uk((size_t)nuniq);
  pbh_dataset_unique_keys(ds, uk.data(), nuniq);
  OK(pbh_dataset_destroy(ds));

  // host table: bulk create / pull / push from the pool, SSD tier, resharded load
  PbxSparseOptConfig cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.nonclk_coeff = 0.1f; cfg.clk_coeff = 1.f; cfg.min_bound = -10; cfg.max_bound = 10; cfg.learning_rate = 0.05f;
  cfg.initial_g2sum = 3; cfg.mf_create_thresholds = 0; cfg.mf_learning_rate = 0.05f; cfg.mf_initial_g2sum = 3;
  cfg.mf_initial_range = 1e-3f; cfg.mf_min_bound = -10; cfg.mf_max_bound = 10;
  void* tab = nullptr;
  OK(pbh_table_create(&tab, 8, 0, &cfg, 1));
  PbxRowLayout L = pbx_make_layout(8, 0);
  std::vector<float> rows(uk.size() * (size_t)L.stride);
  OK(pbh_table_fetch_rows(tab, uk.data(), (int64_t)uk.size(), rows.data(), 1));
  std::vector<float> push(uk.size() * (size_t)L.push_stride, 0.f);
  for (size_t i = 0; i < uk.size(); ++i) { push[i * L.push_stride + 1] = 1.f; push[i * L.push_stride + 3] = 0.1f; }
  OK(pbh_table_push(tab, uk.data(), (int64_t)uk.size(), push.data(), L.push_stride));
  std::vector<float> pulled(uk.size() * 11);
  OK(pbh_table_pull(tab, uk.data(), (int64_t)uk.size(), pulled.data(), 11));
  const std::string ck = dir + "/san-ckpt", ssd = dir + "/san-ssd";
  std::string mk = "mkdir -p " + ck + " " + ssd;
  CHECK(system(mk.c_str()) == 0);
  int64_t nsaved = 0;
  OK(pbh_table_save(tab, ck.c_str(), 1, 0, 0.f, 0.f, 1, &nsaved));        // delta: clears the dirty flags
  OK(pbh_table_ssd_open(tab, ssd.c_str()));
  CHECK(pbh_table_spill(tab, -1.f, (int64_t)uk.size() / 3) > 0);
  OK(pbh_table_pull(tab, uk.data(), (int64_t)uk.size() / 2, pulled.data(), 11));   // promotes half
  OK(pbh_table_save(tab, ck.c_str(), 0, 0, 0.f, 0.f, 0, &nsaved));
  CHECK(nsaved == (int64_t)uk.size());
  CHECK(pbh_table_load_ssd2mem(tab) >= 0 && pbh_table_ssd_rows(tab) == 0);
  void* tab2 = nullptr;
  OK(pbh_table_create(&tab2, 8, 0, &cfg, 2));
  int64_t n0 = 0, total = 0;
  for (int shard = 0; shard < 64; ++shard) {
    char name[600];
    snprintf(name, sizeof(name), "%s/part-000-%05d", ck.c_str(), shard);
    OK(pbh_table_load_file_owned(tab2, name, 0, 1, 3, &n0));
    total += n0;
  }
  CHECK(total > 0 && total < (int64_t)uk.size() && pbh_table_size(tab2) == total);
  // several bulk calls on ONE table at the same time (the pass builder's fetch threads; a write-back next to
  // the next pass's BuildPull): shard locks inside the calls, a shared table lock around them
  {
    std::vector<std::thread> ths;
    std::vector<uint64_t> ck(60000);
    for (size_t i = 0; i < ck.size(); ++i) ck[i] = 7000000ull + (uint64_t)i * 3;
    for (int t = 0; t < 4; ++t)
      ths.emplace_back([&, t] {
        const size_t n = ck.size() / 2, off = (size_t)t * (ck.size() / 8);
        std::vector<float> r(n * (size_t)L.stride);
        pbh_table_fetch_rows(tab, ck.data() + off, (int64_t)n, r.data(), 1);
        if (t % 2) pbh_table_store_rows(tab, ck.data() + off, (int64_t)n, r.data(), 1);
        std::vector<float> pl(n * 11);
        pbh_table_pull(tab, ck.data() + off, (int64_t)n, pl.data(), 11);
      });
    std::atomic<int> bad_size{0};
    ths.emplace_back([&] { for (int i = 0; i < 20; ++i) if (pbh_table_size(tab) <= 0) bad_size++; });
    for (auto& th : ths) th.join();
    CHECK(bad_size.load() == 0);
    std::vector<float> r(ck.size() * (size_t)L.stride);
    OK(pbh_table_fetch_rows(tab, ck.data(), (int64_t)ck.size(), r.data(), 0));
  }
  // damaged model files: bytes knocked out, lines cut, fields removed -- loaded into fresh tables (a refusal
  // or a partial load are both fine; reading out of bounds is not)
  for (int round = 0; round < 16; ++round) {
    char name[600];
    snprintf(name, sizeof(name), "%s/part-000-%05d", ck.c_str(), round % 64);
    std::vector<char> bytes;
    { FILE* f = fopen(name, "rb"); if (!f) continue; int ch; while ((ch = fgetc(f)) != EOF) bytes.push_back((char)ch); fclose(f); }
    if (bytes.size() < 64) continue;
    for (int k = 0; k < 30; ++k) bytes[rnd() % bytes.size()] = round % 3 == 0 ? ' ' : (char)(rnd() % 256);
    if (round % 2) bytes.resize(bytes.size() - 1 - rnd() % 40);
    const std::string bad = dir + "/bad-model-part";
    { FILE* f = fopen(bad.c_str(), "wb"); fwrite(bytes.data(), 1, bytes.size(), f); fclose(f); }
    void* t3 = nullptr;
    OK(pbh_table_create(&t3, 8, 0, &cfg, 2));
    int64_t nl = 0;
    pbh_table_load_file_owned(t3, bad.c_str(), round % 4 == 1, 0, 1, &nl);
    CHECK(pbh_table_size(t3) >= 0);
    OK(pbh_table_destroy(t3));
  }
  OK(pbh_table_ssd_close(tab, 1));
  OK(pbh_table_destroy(tab));
  OK(pbh_table_destroy(tab2));

  // key sort / partition
  std::vector<uint64_t> raw(400000);
  for (auto& k : raw) k = ((uint64_t)rnd() << 20) ^ rnd();
  std::vector<uint64_t> copy = raw;
  int64_t nu = pbh_sort_unique_u64(copy.data(), (int64_t)copy.size());
  CHECK(nu > 0 && nu <= (int64_t)raw.size());
  for (int64_t i = 1; i < nu; ++i) CHECK(copy[i - 1] < copy[i]);
  std::vector<uint64_t> parted((size_t)nu);
  int64_t counts[5];
  OK(pbh_partition_by_owner(copy.data(), nu, 5, parted.data(), counts));
  CHECK(counts[0] + counts[1] + counts[2] + counts[3] + counts[4] == nu);
  // text parser on hostile input: random bytes, truncated numbers, huge counts, embedded NULs, very long lines
  {
    const std::string junk = dir + "/san-junk.txt";
    FILE* fp = fopen(junk.c_str(), "w");
    CHECK(fp != nullptr);
    const char* seeds[] = {"1 1.0 3 0.5 0.25 0.1 1 7 1 8 1 9 1 10 1 11 1 12", "1", "1 ", "999999999999 1", "1 1.0 3 0.5",
                           "1 1.0 3 0.5 0.25 0.1 18446744073709551615 1", "1 1e400 3 nan inf -inf 2 1 2", "-1 -1 -1 -1",
                           "1 1.0 3 0.5 0.25 0.1 1 99999999999999999999999999 1 8 1 9 1 10 1 11 1 12", "\t\t \t"};
    for (int rep = 0; rep < 400; ++rep) {
      std::string line = seeds[rnd() % 10];
      int edits = rnd() % 4;
      for (int e = 0; e < edits && !line.empty(); ++e) {
        size_t pos = rnd() % line.size();
        switch (rnd() % 4) {
          case 0: line[pos] = (char)(rnd() & 0xff); if (line[pos] == '\n') line[pos] = ' '; break;
          case 1: line.erase(pos); break;
          case 2: line.insert(pos, std::string(1 + rnd() % 40, (char)('0' + rnd() % 10))); break;
          default: line.insert(pos, " "); break;
        }
      }
      fwrite(line.data(), 1, line.size(), fp);
      fputc('\n', fp);
    }
    std::string longline(300000, '7');
    fprintf(fp, "1 1.0 3 0.5 0.25 0.1 1 %s 1 8 1 9 1 10 1 11 1 12\n", longline.c_str());
    fclose(fp);
    void* dj = make_dataset(S);
    const char* jf[1] = {junk.c_str()};
    OK(pbh_dataset_set_filelist(dj, 1, jf));
    OK(pbh_dataset_set_int(dj, "thread_num", 2));
    OK(pbh_dataset_set_int(dj, "parse_ins_id", 0));
    OK(pbh_dataset_load_into_memory(dj));
    CHECK(pbh_dataset_num_records(dj) >= 0);
    OK(pbh_dataset_set_int(dj, "parse_ins_id", 1));          // same bytes, now with the id / logkey groups expected
    OK(pbh_dataset_set_int(dj, "parse_logkey", 1));
    OK(pbh_dataset_load_into_memory(dj));
    OK(pbh_dataset_prepare_train(dj, 2, 1, 1, 0));
    OK(pbh_dataset_destroy(dj));
  }
  // graph store: build, threaded sampling and walks
  {
    void* g = nullptr;
    OK(pbh_graph_create(&g));
    std::vector<uint64_t> src, dst;
    for (int i = 0; i < 20000; ++i) { src.push_back(1 + rnd() % 3000); dst.push_back(1 + rnd() % 3000); }
    OK(pbh_graph_add_edges(g, 0, src.data(), dst.data(), (int64_t)src.size(), 0));
    OK(pbh_graph_add_edges(g, 1, src.data(), dst.data(), (int64_t)src.size(), 1));
    OK(pbh_graph_build(g, 0));
    OK(pbh_graph_build(g, 1));
    int64_t nn = pbh_graph_num_nodes(g, 0);
    std::vector<uint64_t> ids((size_t)nn), nb((size_t)nn * 4), walks((size_t)nn * 6);
    std::vector<int64_t> rp((size_t)nn + 1), col((size_t)pbh_graph_num_edges(g, 0));
    std::vector<int32_t> cnt((size_t)nn), len((size_t)nn);
    OK(pbh_graph_export_csr(g, 0, ids.data(), rp.data(), col.data()));
    OK(pbh_graph_sample_neighbors(g, 0, ids.data(), nn, 4, 9, nb.data(), cnt.data()));
    int meta[2] = {0, 1};
    OK(pbh_graph_random_walk(g, meta, 2, ids.data(), nn, 6, 3, walks.data(), len.data()));
    for (int64_t i = 0; i < nn; ++i) CHECK(len[i] >= 1 && len[i] <= 6 && cnt[i] >= 0 && cnt[i] <= 4);
    OK(pbh_graph_destroy(g));
  }
  // graph store, round-2 surface: weighted edges + weighted sampling, node files with feature slots (well-formed
  // and damaged), node weights, partitions and their files
  {
    void* g = nullptr;
    OK(pbh_graph_create(&g));
    std::vector<uint64_t> src, dst;
    std::vector<float> w;
    for (int i = 0; i < 8000; ++i) { src.push_back(1 + rnd() % 900); dst.push_back(1 + rnd() % 900); w.push_back((float)(rnd() % 100) / 10.f); }
    w[3] = 0.f;
    OK(pbh_graph_add_edges_weighted(g, 0, src.data(), dst.data(), w.data(), (int64_t)src.size(), 1));
    OK(pbh_graph_build(g, 0));
    CHECK(pbh_graph_is_weighted(g, 0) == 1);
    int64_t nn = pbh_graph_num_nodes(g, 0), ne = pbh_graph_num_edges(g, 0);
    std::vector<uint64_t> ids((size_t)nn), nb((size_t)nn * 5);
    std::vector<int64_t> rp((size_t)nn + 1), col((size_t)ne);
    std::vector<int32_t> cnt((size_t)nn);
    std::vector<float> we((size_t)ne);
    OK(pbh_graph_export_csr(g, 0, ids.data(), rp.data(), col.data()));
    OK(pbh_graph_export_weights(g, 0, we.data()));
    std::vector<std::thread> ths;
    for (int t = 0; t < 3; ++t)
      ths.emplace_back([&, t] {
        std::vector<uint64_t> o((size_t)nn * 5);
        std::vector<int32_t> c((size_t)nn);
        pbh_graph_sample_neighbors_weighted(g, 0, ids.data(), nn, 5, 11 + (uint64_t)t, o.data(), c.data());
      });
    for (auto& th : ths) th.join();
    OK(pbh_graph_sample_neighbors_weighted(g, 0, ids.data(), nn, 5, 3, nb.data(), cnt.data()));
    for (int64_t i = 0; i < nn; ++i) CHECK(cnt[i] >= 0 && cnt[i] <= 5);
    uint64_t unknown[3] = {0, 999999999ull, ids[0]};
    int32_t c3[3];
    uint64_t o3[15];
    OK(pbh_graph_sample_neighbors_weighted(g, 0, unknown, 3, 5, 1, o3, c3));
    CHECK(c3[0] == 0 && c3[1] == 0);
    // weighted edge file and node file, then the same files with bytes knocked out
    std::string ef = dir + "/edges_w.txt", nf = dir + "/nodes.txt", wf = dir + "/node_w.txt";
    {
      FILE* f = fopen(ef.c_str(), "w");
      for (int i = 0; i < 500; ++i) fprintf(f, "%d\t%d\t%.3f\n", 1 + (int)(rnd() % 90), 1 + (int)(rnd() % 90), (double)(rnd() % 50) / 7.0);
      fputs("12\t\n\t\t\nnot a number\t4\t1.0\n77 78\n", f);
      fclose(f);
      f = fopen(nf.c_str(), "w");
      for (int i = 1; i <= 90; ++i) fprintf(f, "user\t%d\t1:%d\t2:%d,%d\t7:%d\n", i, i * 3, i + 1000, i + 2000, i % 5);
      fputs("item\t5\t1:9\nuser\t\nuser\t91\t:\t3:\t:4\tx:y\nuser\t92\t1:18446744073709551615\n", f);
      fclose(f);
      f = fopen(wf.c_str(), "w");
      for (int i = 1; i <= 90; i += 2) fprintf(f, "%d\t%.2f\n", i, i / 10.0);
      fputs("bad\n93\n\t\n", f);
      fclose(f);
    }
    int64_t nl = 0;
    OK(pbh_graph_load_weighted_edge_file(g, 1, ef.c_str(), 0, &nl));
    CHECK(nl >= 500);
    OK(pbh_graph_build(g, 1));
    OK(pbh_graph_load_node_file(g, 0, nf.c_str(), "user", ':', &nl));
    CHECK(nl >= 90);
    int64_t ntn = pbh_graph_num_typed_nodes(g, 0);
    std::vector<uint64_t> tn((size_t)ntn);
    OK(pbh_graph_export_typed_nodes(g, 0, tn.data()));
    int32_t slots[3] = {1, 2, 9};
    std::vector<uint64_t> feat((size_t)ntn * 3);
    OK(pbh_graph_node_features(g, 0, tn.data(), ntn, slots, 3, feat.data()));
    int64_t nf_ids = 0;
    OK(pbh_graph_all_feature_ids(g, 0, nullptr, &nf_ids));
    std::vector<uint64_t> fids((size_t)nf_ids + 1);
    OK(pbh_graph_all_feature_ids(g, 0, fids.data(), &nf_ids));
    OK(pbh_graph_load_node_weight(g, 0, wf.c_str(), &nl));
    std::vector<float> nw((size_t)ntn);
    OK(pbh_graph_node_weights(g, 0, tn.data(), ntn, nw.data()));
    for (int round = 0; round < 12; ++round) {
      std::string bad = dir + "/bad_graph.txt";
      std::vector<char> bytes;
      { FILE* f = fopen(round % 2 ? nf.c_str() : ef.c_str(), "rb"); int ch; while ((ch = fgetc(f)) != EOF) bytes.push_back((char)ch); fclose(f); }
      for (int k = 0; k < 40; ++k) bytes[rnd() % bytes.size()] = (char)(rnd() % 256);
      bytes.resize(bytes.size() - rnd() % 17);
      { FILE* f = fopen(bad.c_str(), "wb"); fwrite(bytes.data(), 1, bytes.size(), f); fclose(f); }
      void* g2 = nullptr;
      OK(pbh_graph_create(&g2));
      int64_t n2 = 0;
      if (round % 2) pbh_graph_load_node_file(g2, 0, bad.c_str(), "user", ':', &n2);
      else { pbh_graph_load_weighted_edge_file(g2, 0, bad.c_str(), 1, &n2); pbh_graph_build(g2, 0); }
      OK(pbh_graph_destroy(g2));
    }
    // partitions: tiny byte budget -> many parts; shapes, exports and files
    int nparts = 0;
    OK(pbh_graph_make_partitions(g, 0, 4096, 2, &nparts));
    CHECK(nparts >= 2 && pbh_graph_partition_num(g, 0) == nparts);
    int64_t tot_nodes = 0, tot_edges = 0;
    for (int p = 0; p < nparts; ++p) {
      int64_t pn = 0, pe = 0;
      OK(pbh_graph_partition_shape(g, 0, p, &pn, &pe));
      std::vector<uint64_t> pid((size_t)pn), pcol((size_t)pe);
      std::vector<int64_t> prp((size_t)pn + 1);
      std::vector<float> pw((size_t)pe);
      OK(pbh_graph_export_partition(g, 0, p, pid.data(), prp.data(), pcol.data(), pw.data()));
      CHECK(prp[(size_t)pn] == pe);
      tot_nodes += pn; tot_edges += pe;
    }
    CHECK(tot_nodes == nn && tot_edges == ne);
    for (int i = 0; i < nparts + 2; ++i) pbh_graph_next_partition(g, 0);
    OK(pbh_graph_export_partition_files(g, 0, dir.c_str()));
    CHECK(pbh_graph_partition_shape(g, 0, nparts + 5, &tot_nodes, &tot_edges) != 0);
    OK(pbh_graph_destroy(g));
  }
  // metrics: concurrent adds, WuAUC, state export / import
  {
    void* a = nullptr;
    OK(pbh_auc_create(&a, 10000));
    std::vector<std::thread> ths;
    for (int t = 0; t < 4; ++t)
      ths.emplace_back([&, t] {
        std::vector<float> p(5000), y(5000);
        std::vector<uint64_t> uid(5000);
        unsigned r = 77u + (unsigned)t;
        for (size_t i = 0; i < p.size(); ++i) { r = r * 1664525u + 1013904223u; p[i] = (r >> 8) / 16777216.f; y[i] = (r & 64) ? 1.f : 0.f; uid[i] = r % 50; }
        pbh_auc_add(a, p.data(), y.data(), nullptr, (int64_t)p.size(), 0);
        pbh_auc_add_uid(a, p.data(), y.data(), uid.data(), (int64_t)p.size());
      });
    for (auto& th : ths) th.join();
    double out8[8], out4[4];
    std::vector<double> tab(2 * 10000), sums(5);
    OK(pbh_auc_compute(a, out8));
    OK(pbh_auc_compute_wuauc(a, out4));
    CHECK(out8[7] == 20000 && out8[0] > 0.3 && out8[0] < 0.7);
    OK(pbh_auc_export(a, tab.data(), sums.data()));
    OK(pbh_auc_import(a, tab.data(), sums.data()));
    OK(pbh_auc_destroy(a));
  }
  // async dense table: pushes from two threads while a third pulls
  {
    void* ad = nullptr;
    OK(pbh_async_dense_create(&ad, 4096, 0.01f, 3, 4));
    std::vector<float> w(4096, 0.5f);
    OK(pbh_async_dense_set_params(ad, w.data()));
    OK(pbh_async_dense_set_lr_range(ad, 0, 100, 0.f));
    std::vector<std::thread> ths;
    for (int t = 0; t < 2; ++t)
      ths.emplace_back([&] { std::vector<float> g(4096, 0.1f); for (int i = 0; i < 20; ++i) pbh_async_dense_push(ad, g.data()); });
    ths.emplace_back([&] { std::vector<float> o(4096); for (int i = 0; i < 20; ++i) pbh_async_dense_pull(ad, o.data()); });
    for (auto& th : ths) th.join();
    OK(pbh_async_dense_wait(ad));
    OK(pbh_async_dense_pull(ad, w.data()));
    CHECK(w[0] == 0.5f && w[200] < 0.5f && pbh_async_dense_updates(ad) > 0);
    OK(pbh_async_dense_destroy(ad));
  }
  // input index from several threads
  {
    void* ix = nullptr;
    OK(pbh_input_index_create(&ix));
    std::vector<std::thread> ths;
    for (int t = 0; t < 4; ++t)
      ths.emplace_back([&, t] {
        for (int i = 0; i < 2000; ++i) {
          std::string k = "key" + std::to_string((i * 7 + t) % 500);
          int fresh = 0;
          pbh_input_index_add(ix, k.c_str(), k.size(), &fresh);
          pbh_input_index_get(ix, k.c_str(), k.size());
        }
      });
    for (auto& th : ths) th.join();
    CHECK(pbh_input_index_size(ix) == 500 && pbh_input_index_miss(ix) == 0);
    OK(pbh_input_index_destroy(ix));
  }
  printf("sanitize driver ok: %lld records, %lld unique keys\n", (long long)FILES * LINES, (long long)nuniq);
  return 0;
}
