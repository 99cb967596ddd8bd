/*
 * sam2bam — TEST / BENCH INFRASTRUCTURE (never linked into the product).
 *
 * Turns the coordinate-sorted SAM text that margin_b200/synth.py writes for the synthetic polish workloads
 * ("BAM mode") into the indexed BAM `margin polish` reads through htsIntegration.c
 * (bamChunker_construct2, impl/htsIntegration.c:203; convertToReadsAndAlignmentsWithFiltered, :557).  There is no
 * samtools / pysam on the box, so this links the reference's vendored htslib 1.9 objects built by oracle/Makefile.
 *
 * usage: sam2bam in.sam out.bam      (writes out.bam and out.bam.bai)
 */
#include <stdio.h>
#include <stdlib.h>
#include "htslib/sam.h"

int main(int argc, char **argv) {
    if (argc != 3) {
        fprintf(stderr, "usage: %s in.sam out.bam\n", argv[0]);
        return 2;
    }
    samFile *in = sam_open(argv[1], "r");
    if (in == NULL) {
        fprintf(stderr, "sam2bam: cannot open %s\n", argv[1]);
        return 1;
    }
    bam_hdr_t *hdr = sam_hdr_read(in);
    samFile *out = sam_open(argv[2], "wb");
    if (hdr == NULL || out == NULL || sam_hdr_write(out, hdr) != 0) {
        fprintf(stderr, "sam2bam: cannot start %s\n", argv[2]);
        return 1;
    }
    bam1_t *rec = bam_init1();
    long n = 0;
    int r;
    while ((r = sam_read1(in, hdr, rec)) >= 0) {
        if (sam_write1(out, hdr, rec) < 0) {
            fprintf(stderr, "sam2bam: write failed at record %ld\n", n);
            return 1;
        }
        n++;
    }
    if (r < -1) {
        fprintf(stderr, "sam2bam: parse error after record %ld\n", n);
        return 1;
    }
    bam_destroy1(rec);
    bam_hdr_destroy(hdr);
    sam_close(in);
    if (sam_close(out) != 0) return 1;
    if (sam_index_build(argv[2], 0) != 0) {
        fprintf(stderr, "sam2bam: index build failed (records must be coordinate sorted)\n");
        return 1;
    }
    fprintf(stderr, "sam2bam: %ld records -> %s\n", n, argv[2]);
    return 0;
}
