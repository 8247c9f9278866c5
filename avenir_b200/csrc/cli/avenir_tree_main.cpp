// avenir-tree: command-line host with the reference's job syntax
//   hadoop jar avenir-1.0.jar org.avenir.tree.DecisionTreeBuilder -Dconf.path=detr.properties IN OUT
// becomes
//   avenir-tree org.avenir.tree.DecisionTreeBuilder -Dconf.path=detr.properties IN OUT
// (resource/detr.sh:52; resource/abandoned_shopping_cart_retarget_tutorial.txt:29-39).  Exit 0/1.
#include <cstdio>

#include "../../../include/avenir_tree.h"

int main(int argc, char** argv) {
    if (argc < 2) {
        fprintf(stderr, "usage: %s <DecisionTreeBuilder|ClassPartitionGenerator> -Dconf.path=<file.properties> [-Dkey=value ...] <inPath> <outPath>\n", argv[0]);
        return 2;
    }
    char err[2048];
    int rc = avt_tool_run(argv[1], argc - 2, (const char* const*)(argv + 2), err, sizeof err);
    if (rc != 0) fprintf(stderr, "%s failed: %s\n", argv[1], err);
    return rc;
}
