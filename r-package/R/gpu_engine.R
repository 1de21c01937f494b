## Drop-in bodies for DiceOptim's criteria on top of the .Call shim (r-package/src/dk_rcall.c).
## Signatures are the reference's (R/EI.R:80, R/EI.grad.R:97, R/qEI.R:119, R/qEI.grad.R:267, R/AKG.R:93,
## R/kriging.quantile.R:92); only the bodies change.  Untested here: the build image has no R.

.dk_cache <- new.env(parent = emptyenv())

dk_handle <- function(model, device = 0L) {
  key <- paste0(digest_T(model@T), "@", device)
  h <- .dk_cache[[key]]
  if (is.null(h)) {
    cov <- model@covariance
    kern <- match(cov@name, c("gauss", "exp", "matern3_2", "matern5_2", "powexp")) - 1L
    if (is.na(kern)) stop("covariance kernel not supported by the GPU engine: ", cov@name)
    trend <- if (model@trend.formula == ~1) 0L else if (ncol(model@F) == model@d + 1L) 1L else
      stop("only trend formulas ~1 and ~. are supported by the GPU engine")
    slots <- list(T = model@T, z = as.numeric(model@z), M = as.matrix(model@M), X = as.matrix(model@X),
                  y = as.numeric(model@y), F = as.matrix(model@F), trend.coef = as.numeric(model@trend.coef),
                  range.val = as.numeric(cov@range.val),
                  shape.val = if (cov@name == "powexp") as.numeric(cov@shape.val) else numeric(0),
                  noise.var = if (model@noise.flag) as.numeric(model@noise.var) else numeric(0),
                  n = model@n, d = model@d, p = ncol(model@F), kernel = kern, trend = trend, sd2 = cov@sd2,
                  nugget = if (cov@nugget.flag) cov@nugget else 0, nugget.flag = cov@nugget.flag)
    h <- .Call("dkR_model_create", slots, as.integer(device))
    .dk_cache[[key]] <- h
  }
  h
}
digest_T <- function(T) paste(dim(T)[1], format(sum(T), digits = 17), format(T[length(T)], digits = 17))
.type_id <- function(type) c(SK = 0L, UK = 1L)[[type]]
.as_rows <- function(x, d) { if (is.null(dim(x))) x <- matrix(as.numeric(x), nrow = 1)
  if (ncol(x) != d) stop("x does not have the right size"); storage.mode(x) <- "double"; as.matrix(x) }

EI <- function(x, model, plugin = NULL, type = "UK", minimization = TRUE, envir = NULL, proxy = FALSE) {
  if (is.null(plugin)) plugin <- if (minimization) min(model@y) else -max(model@y)
  x <- .as_rows(x, model@d)
  r <- .Call("dkR_ei", dk_handle(model), x, plugin, .type_id(type), minimization, proxy, !is.null(envir))
  if (!is.null(envir)) {
    km <- if (minimization) r[[2]] else -r[[2]]
    xcr <- if (proxy) NULL else (plugin - km) / r[[3]]
    assign("xcr", xcr, envir = envir)
    assign("xcr.prob", if (proxy) NULL else pnorm(xcr), envir = envir)
    assign("xcr.dens", if (proxy) NULL else dnorm(xcr), envir = envir)
    assign("kriging.sd", r[[3]], envir = envir)
  }
  r[[1]]
}
EI.grad <- function(x, model, plugin = NULL, type = "UK", minimization = TRUE, envir = NULL, proxy = FALSE) {
  if (is.null(plugin)) plugin <- if (minimization) min(model@y) else -max(model@y)
  g <- .Call("dkR_ei_grad", dk_handle(model), .as_rows(x, model@d), plugin, .type_id(type), minimization, proxy)
  if (is.null(dim(x))) as.numeric(g) else g
}
kriging.quantile <- function(x, model, beta = 0.1, type = "UK", envir = NULL)
  .Call("dkR_quantile", dk_handle(model), .as_rows(x, model@d), qnorm(beta), .type_id(type), FALSE)[[1]]
kriging.quantile.grad <- function(x, model, beta = 0.1, type = "UK", envir = NULL)
  as.numeric(.Call("dkR_quantile", dk_handle(model), .as_rows(x, model@d), qnorm(beta), .type_id(type), TRUE)[[2]])
AKG <- function(x, model, new.noise.var = 0, type = "UK", envir = NULL)
  .Call("dkR_akg", dk_handle(model), .as_rows(x, model@d), new.noise.var, .type_id(type))
qEI <- function(x, model, plugin = NULL, type = "UK", minimization = TRUE, fastCompute = TRUE, eps = 1e-5, envir = NULL) {
  if (!is.matrix(x)) x <- matrix(x, ncol = model@d)
  if (is.null(plugin)) plugin <- if (minimization) min(model@y) else max(model@y)
  .Call("dkR_qei", dk_handle(model), x, nrow(x), plugin, .type_id(type), minimization, FALSE)
}
qEI.grad <- function(x, model, plugin = NULL, type = "UK", minimization = TRUE, fastCompute = TRUE, eps = 1e-6, envir = NULL) {
  if (!is.matrix(x)) x <- matrix(x, ncol = model@d)
  if (is.null(plugin)) plugin <- if (minimization) min(model@y) else -max(model@y)
  if (nrow(x) == 1) return(matrix(EI.grad(x, model, plugin, type), nrow = 1))
  .Call("dkR_qei", dk_handle(model), x, nrow(x), plugin, .type_id(type), minimization, TRUE)
}
## the batched screen of max_crit (R/max_crit.R:104-110) in one call: list(index (1-based), value)
screen_EI <- function(x, model, plugin = NULL, type = "UK", minimization = TRUE, proxy = FALSE, k = 1L) {
  if (is.null(plugin)) plugin <- if (minimization) min(model@y) else -max(model@y)
  .Call("dkR_screen_ei", dk_handle(model), .as_rows(x, model@d), plugin, .type_id(type), minimization, proxy, as.integer(k))
}
## checkPredict (R/checkPredict.R:18-75), vectorised over the rows of x; fastfun members of `model` are skipped as there
checkPredict <- function(x, model, threshold = 1e-4, distance = "covdist", type = "UK") {
  if (is.null(distance)) distance <- "covdist"
  if (is.null(threshold)) threshold <- 1e-4
  kms <- Filter(function(m) is(m, "km"), model)
  x <- .as_rows(x, kms[[1]]@d)
  if (distance == "none") return(rep(FALSE, nrow(x)))
  id <- match(distance, c("euclidean", "covdist", "covratio")) - 1L
  .Call("dkR_check_predict", lapply(kms, dk_handle), x, .type_id(type), id) < threshold
}
## crit_EFI (R/crit_EFI.R:125-228) for every row of x: one upload of x, EI and the Feas factors multiplied on the device
crit_EFI <- function(x, model.fun, model.constraint, equality = FALSE, critcontrol = NULL, plugin = NULL, type = "UK") {
  if (!is(model.constraint, "list")) model.constraint <- list(model.constraint)
  n.cst <- length(model.constraint)
  if (is.null(plugin)) {
    observations <- sapply(model.constraint, function(m) m@y)
    feasibility <- test_feas_vec(cst = observations, equality = equality, tolConstraints = critcontrol$tolConstraints)
    plugin <- if (any(feasibility)) min(model.fun@y[feasibility]) else Inf
  }
  equalityvec <- if (length(equality) == 1 && equality == FALSE) rep(FALSE, n.cst) else equality
  tolvec <- critcontrol$tolConstraints
  if (is.null(tolvec)) { tolvec <- rep(0, n.cst); tolvec[equalityvec] <- 0.05 }
  x <- .as_rows(x, model.fun@d)
  out <- .Call("dkR_feasible_improvement", dk_handle(model.fun), lapply(model.constraint, dk_handle), x, plugin,
               .type_id(type), 0L, as.logical(equalityvec), as.double(tolvec))
  out[checkPredict(x, c(model.constraint, model.fun), type = type, distance = critcontrol$distance,
                   threshold = critcontrol$threshold)] <- 0
  out
}
