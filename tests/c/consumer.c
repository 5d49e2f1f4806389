/* A C consumer of the drop-in: includes the header as plain C, links the
 * shared library and calls the reference's own entry points (void return,
 * struct pf_bitmap in and out), exactly as a program written against
 * libpillowfight would.  Usage: consumer <op> <width> <height> <in.raw> <out.raw>
 * (op: sobel | gaussian | canny | ace; raw RGBA8 files).  Exit 0 on success,
 * 2 when the library reports an error. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "pillowfight_b200.h"

int main(int argc, char **argv)
{
	struct pf_bitmap in, out;
	size_t n;
	FILE *f;

	if (argc != 6)
		return 64;
	in.size.x = out.size.x = atoi(argv[2]);
	in.size.y = out.size.y = atoi(argv[3]);
	n = (size_t)in.size.x * in.size.y;
	in.pixels = malloc(n * sizeof(union pf_pixel));
	out.pixels = malloc(n * sizeof(union pf_pixel));
	if (!in.pixels || !out.pixels)
		return 65;
	f = fopen(argv[4], "rb");
	if (!f || fread(in.pixels, sizeof(union pf_pixel), n, f) != n)
		return 66;
	fclose(f);

	if (!strcmp(argv[1], "sobel")) {
		pf_apply_sobel(&in, &out);
	} else if (!strcmp(argv[1], "gaussian")) {
		pf_apply_gaussian(&in, &out, 2.0, 5);
	} else if (!strcmp(argv[1], "canny")) {
		pf_apply_canny(&in, &out);
	} else if (!strcmp(argv[1], "ace")) {
		srand(12345);                     /* as the reference's Python glue does before pf_ace */
		pf_apply_ace(&in, &out, 40, 10.0, 1000.0, 4);
	} else {
		return 64;
	}
	if (pf_b200_last_error() != PF_B200_OK) {
		fprintf(stderr, "consumer: %s\n", pf_b200_last_error_string());
		return 2;
	}
	f = fopen(argv[5], "wb");
	if (!f || fwrite(out.pixels, sizeof(union pf_pixel), n, f) != n)
		return 67;
	fclose(f);
	free(in.pixels);
	free(out.pixels);
	return 0;
}
