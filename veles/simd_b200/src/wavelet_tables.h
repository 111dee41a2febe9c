/* wavelet_tables.h -- lowpass filter tables of the wavelet path (generated data in
 * wavelet_tables.c; see tools/gen_wavelet_tables.py).  Row k of the Daubechies and Symlet
 * tables holds the 2(k+1)-tap filter, row k of the Coiflet table the 6(k+1)-tap filter, as the
 * reference indexes kDaubechiesF / kSymletsF / kCoifletsF (/root/reference/src/wavelet.c:
 * 190-206). */
#ifndef B200_WAVELET_TABLES_H_
#define B200_WAVELET_TABLES_H_

extern const float b200_daubechies[38][76];
extern const float b200_symlets[38][76];
extern const float b200_coiflets[5][30];

#endif
