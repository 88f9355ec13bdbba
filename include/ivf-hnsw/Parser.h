// <ivf-hnsw/Parser.h> of the B200 drop-in: same option names, member names and defaults-by-omission as
// the reference's command-line parser (reference Parser.h:14-190), table driven.
#pragma once
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>

struct Parser {
    const char *cmd;
    // HNSW parameters (accepted for compatibility; the B200 coarse quantizer is exact)
    size_t M, efConstruction;
    // data
    size_t nb, nt, nsubt, nc, nsubc, nq, ngt, d;
    // PQ
    size_t code_size;
    bool do_opq;
    // search
    size_t k, nprobe, max_codes, efSearch;
    bool do_pruning;
    // paths
    const char *path_base, *path_learn, *path_q, *path_gt, *path_centroids, *path_precomputed_idxs, *path_info,
        *path_edges, *path_pq, *path_opq_matrix, *path_norm_pq, *path_index;

    Parser(int argc, char **argv)
    {
        memset(static_cast<void *>(this), 0, sizeof(*this));
        cmd = argv[0];
        if (argc == 1) usage();
        struct SizeOpt { const char *name; size_t Parser::*field; };
        struct PathOpt { const char *name; const char *Parser::*field; };
        struct FlagOpt { const char *name; bool Parser::*field; };
        static const SizeOpt sizes[] = {
            {"-M", &Parser::M}, {"-efConstruction", &Parser::efConstruction}, {"-nb", &Parser::nb},
            {"-nc", &Parser::nc}, {"-nsubc", &Parser::nsubc}, {"-nt", &Parser::nt}, {"-nsubt", &Parser::nsubt},
            {"-nq", &Parser::nq}, {"-ngt", &Parser::ngt}, {"-d", &Parser::d}, {"-code_size", &Parser::code_size},
            {"-k", &Parser::k}, {"-nprobe", &Parser::nprobe}, {"-max_codes", &Parser::max_codes},
            {"-efSearch", &Parser::efSearch}};
        static const PathOpt paths[] = {
            {"-path_base", &Parser::path_base}, {"-path_learn", &Parser::path_learn}, {"-path_q", &Parser::path_q},
            {"-path_gt", &Parser::path_gt}, {"-path_centroids", &Parser::path_centroids},
            {"-path_precomputed_idx", &Parser::path_precomputed_idxs}, {"-path_info", &Parser::path_info},
            {"-path_edges", &Parser::path_edges}, {"-path_pq", &Parser::path_pq},
            {"-path_opq_matrix", &Parser::path_opq_matrix}, {"-path_norm_pq", &Parser::path_norm_pq},
            {"-path_index", &Parser::path_index}};
        static const FlagOpt flags[] = {{"-opq", &Parser::do_opq}, {"-pruning", &Parser::do_pruning}};
        for (int i = 1; i < argc; i++) {
            const char *a = argv[i];
            if (!strcmp(a, "-h") || !strcmp(a, "--help")) usage();
            if (i == argc - 1) break;
            bool known = false;
            for (const SizeOpt &o : sizes)
                if (!known && !strcmp(a, o.name)) { sscanf(argv[++i], "%zu", &(this->*o.field)); known = true; }
            for (const PathOpt &o : paths)
                if (!known && !strcmp(a, o.name)) { this->*o.field = argv[++i]; known = true; }
            for (const FlagOpt &o : flags)
                if (!known && !strcmp(a, o.name)) { this->*o.field = !strcmp(argv[++i], "on"); known = true; }
        }
    }

    void usage()
    {
        printf("Usage: %s [options]\n", cmd);
        printf("  HNSW (ignored, exact coarse search): -M -efConstruction -efSearch\n");
        printf("  data: -nb -nc -nsubc -nt -nsubt -nq -ngt -d\n");
        printf("  PQ: -code_size <bytes> -opq on|off\n");
        printf("  search: -k -nprobe -max_codes -pruning on|off\n");
        printf("  paths: -path_base -path_learn -path_q -path_gt -path_centroids -path_precomputed_idx\n");
        printf("         -path_info -path_edges -path_pq -path_opq_matrix -path_norm_pq -path_index\n");
        exit(0);
    }
};
