/* TEST INFRASTRUCTURE: empty stand-in; the mock types come from oracle/ref_mock.h and the probe */
