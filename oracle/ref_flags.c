/* TEST INFRASTRUCTURE — not part of the product.
 *
 * Writable definitions of the reference's nine link-time model flags
 * (`extern const bool GP;` ... /root/reference/src/RVmodel.h:16-34 and
 * `extern const bool hyperpriors;` src/RVConditionalPrior.h:11).  A user's
 * kima_setup.cpp defines them as constants; defining them here as plain C
 * globals of the same unmangled names lets one libkima_ref.so serve every flag
 * combination.  They are only changed while no RVmodel is alive. */
#include <stdbool.h>

bool GP = false;
bool MA = false;
bool hyperpriors = false;
bool trend = false;
int degree = 0;
bool multi_instrument = false;
bool known_object = false;
int n_known_object = 0;
bool studentt = false;

void ref_flags_store(int GP_, int MA_, int hyper_, int trend_, int degree_,
                     int multi_, int ko_, int nko_, int studentt_)
{
    GP = GP_ != 0;
    MA = MA_ != 0;
    hyperpriors = hyper_ != 0;
    trend = trend_ != 0;
    degree = degree_;
    multi_instrument = multi_ != 0;
    known_object = ko_ != 0;
    n_known_object = nko_;
    studentt = studentt_ != 0;
}
