# R-facing drop-in for PPLS::PPLS_simult (Package/PPLS/R/EM_W_multi.R:758-807) on the B200 engine.
#
# Same signature, same returned list (Expectations / loglik / estimates, class "PPLS_simult").
# What stays in R, as in the reference: argument handling (:759-761), the random start through the
# reference's own sequential fit PPLS(X, Y, a, 20, 1e-4, 'random') with its three attempts (:762-771),
# and both sign/order normalisations (:773-778, :794-799), which act on r-sized objects.
# What moves to the GPU: the loop :780-793 and the final E-step :802, one .Call into the Rcpp shim
# r/ppls_b200_shim.cpp, which binds the C ABI include/ppls_b200.h (ppls_simult_run).
#
# NOT BUILT OR RUN IN THIS REPOSITORY'S CONTAINERS (R, Rcpp are not installed there): the same C ABI is
# exercised through ctypes by ppls_b200/api.py, whose PPLS_simult mirrors this file line for line.

.ppls_b200_normalise <- function(W, C, B, sigT, a) {
  signLoad <- sign(diag(sigT %*% B))
  rotLoad <- order(diag(sigT %*% B %*% diag(signLoad, a)), decreasing = TRUE)
  list(W = W[, rotLoad, drop = FALSE] %*% diag(signLoad, a),
       C = C[, rotLoad, drop = FALSE] %*% diag(signLoad, a),
       B = diag(diag(B %*% diag(signLoad, a))[rotLoad], a),
       sigT = diag(diag(sigT)[rotLoad], a))
}

# Upload X and Y once and keep them on the GPU(s): the returned handle (an external pointer) can be passed to any number
# of PPLS_simult_b200 calls (`handle =`), which then skip the host-to-device copy -- 86 % of a 20-step call at
# N = 1e7, p = 1000, q = 200.  device = 0:7 splits the rows over eight GPUs of one node (ppls_ctx_create_multi: this one R
# process drives them all).  Freed by ppls_b200_close() or when R collects the handle.
ppls_b200_open <- function(X, Y, device = 0L) {
  stopifnot(is.matrix(X), is.matrix(Y), nrow(X) == nrow(Y))
  storage.mode(X) <- "double"; storage.mode(Y) <- "double"
  .Call("ppls_b200_open", X, Y, as.integer(device), PACKAGE = "pplsb200")
}
ppls_b200_close <- function(handle) invisible(.Call("ppls_b200_close", handle, PACKAGE = "pplsb200"))

PPLS_simult_b200 <- function(X, Y, a, EMsteps = 10, atol = 1e-4, type = c("SVD", "QR"), ...,
                             init = NULL, device = 0L, handle = NULL) {
  type <- match.arg(type)
  dots <- list(...)
  if (isTRUE(dots$debug))
    warning("debug=TRUE (dense Expect_M) is not part of the GPU path; using the fast E-step")
  stopifnot(is.matrix(X), is.matrix(Y), nrow(X) == nrow(Y), a >= 1, ncol(X) >= a, ncol(Y) >= a)
  storage.mode(X) <- "double"; storage.mode(Y) <- "double"
  if (is.null(init)) {
    f0 <- try(suppressWarnings(PPLS::PPLS(X, Y, a, 20, 1e-4, "random")), TRUE)
    if (inherits(f0, "try-error")) f0 <- try(suppressWarnings(PPLS::PPLS(X, Y, a, 20, 1e-4, "random")), TRUE)
    if (inherits(f0, "try-error")) f0 <- try(suppressWarnings(PPLS::PPLS(X, Y, a, 20, 1e-4, "random")), TRUE)
    init <- list(W = f0$W, C = f0$C, B = diag(f0$B, a), sigE = f0$sig[a, 1], sigF = f0$sig[a, 2],
                 sigH = f0$sig[a, 3], sigT = diag(f0$sig[, 4], a))
  }
  st <- .ppls_b200_normalise(init$W, init$C, init$B, init$sigT, a)
  own <- is.null(handle)
  if (own) { handle <- ppls_b200_open(X, Y, device); on.exit(ppls_b200_close(handle), add = TRUE) }
  res <- .Call("PPLS_simult_b200_run", handle, st$W, st$C, diag(st$B), init$sigE, init$sigF, init$sigH,
               diag(st$sigT), as.integer(EMsteps), as.numeric(atol), type, PACKAGE = "pplsb200")
  fin <- .ppls_b200_normalise(res$W, res$C, diag(res$B, a), diag(res$sigT, a), a)
  outp <- list(W = fin$W, C = fin$C, B = fin$B, sigE = res$sigE, sigF = res$sigF, sigH = res$sigH,
               sigT = fin$sigT)
  logl_incr <- res$loglik
  if (any(diff(logl_incr) < 0)) warning("Negative increments of likelihood")
  I <- diag(1, a)
  Eout <- list(mu_T = res$mu_T, mu_U = res$mu_U, Ctt = diag(res$Ctt, a), Cuu = diag(res$Cuu, a),
               Cut = diag(res$Cut, a), Cee = matrix(res$Cee, 1, 1), Cff = matrix(res$Cff, 1, 1),
               Chh = res$Chh)
  outpt <- list(Expectations = Eout, loglik = logl_incr, estimates = outp)
  class(outpt) <- "PPLS_simult"
  outpt
}
