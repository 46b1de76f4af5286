;;;; im2col.lisp — Im2ColNode / Col2ImNode on CUDATensor.
;;;; Counterpart of `source/backends/cpu/im2col.lisp:4-50` (which calls `waffe2_{im2col,col2im}_s_column`,
;;;; `cl-waffe2-simd/api.lisp:27-61`).  The node definitions (`source/base-impl/unfold.lisp:9-110`), `unfold`,
;;;; `Conv2D`, `MaxPool2D`, `AvgPool2D` (`source/nn/{im2col,conv,pool}.lisp`) stay as they are: they are
;;;; compositions of nodes and run unchanged on this device.
(in-package :cl-waffe2/backends.cuda)

(defun unfold-reject-p (&rest args)
  "Decline (fall through to the next device) for what libwaffe2cuda does not cover: non-float dtypes are
rejected by `cuda-reject-p`; `:order :row` is not offered because the reference's `_row` kernels index COL
with strides that leave the buffer (`cl-waffe2-simd/kernels/im2col.c:245-256`)."
  (apply #'cuda-reject-p args))

(defmacro with-unfold-geom ((var node x-shape) &body body)
  "Fill a foreign wf2cu_unfold_geom from the slots of an Im2ColNode / Col2ImNode and the (N C H W) image shape."
  `(with-foreign-object (,var '(:struct wf2cu-unfold-geom))
     (with-slots (N C h-out w-out k-h k-w stride-h stride-w padding-h padding-w dilation-h dilation-w) ,node
       (macrolet ((put (slot value)
		    `(setf (foreign-slot-value ,',var '(:struct wf2cu-unfold-geom) ',slot) (observe ,value))))
	 (put n (nth 0 ,x-shape)) (put c C) (put h (nth 2 ,x-shape)) (put w (nth 3 ,x-shape))
	 (put h-out h-out) (put w-out w-out) (put k-h k-h) (put k-w k-w)
	 (put pad-h padding-h) (put pad-w padding-w) (put stride-h stride-h) (put stride-w stride-w)
	 (put dilation-h dilation-h) (put dilation-w dilation-w)))
     ,@body))

(define-impl-op (Im2ColNode :device CUDATensor :reject-p #'unfold-reject-p)
  :forward ((self x col)
	    ;; x is the padded image, col the (N C k-h k-w h-out w-out) buffer; both dense, as in the CPU impl
	    (setf (h-of self) (nth 2 (shape x))
		  (w-of self) (nth 3 (shape x)))
	    (unless (eql (order x) :column)
	      (error "CUDATensor Im2ColNode: only :order :column is supported"))
	    (with-unfold-geom (g self (shape x))
	      (im2col (device-vec-ptr (tensor-vec col)) (tensor-initial-offset col)
		      (device-vec-ptr (tensor-vec x)) (tensor-initial-offset x) g))
	    col))

(define-impl-op (Col2ImNode :device CUDATensor :reject-p #'unfold-reject-p)
  :forward ((self col output-to)
	    (setf (h-of self) (nth 2 (shape output-to))
		  (w-of self) (nth 3 (shape output-to)))
	    (unless (eql (order col) :column)
	      (error "CUDATensor Col2ImNode: only :order :column is supported"))
	    ;; the incoming gradient is a permuted view (Permute-Node backward); the kernel wants it dense
	    (let ((col* (if (cl-waffe2/base-impl::tensor-permuted-p col) (cuda-contiguous-copy col) col)))
	      (with-unfold-geom (g self (shape output-to))
		(col2im (device-vec-ptr (tensor-vec col*)) (tensor-initial-offset col*)
			(device-vec-ptr (tensor-vec output-to)) (tensor-initial-offset output-to) g)))
	    output-to))

(defun cuda-contiguous-copy (tensor)
  "A dense copy of a permuted / viewed CUDATensor (what `->contiguous` builds as a graph node), run now."
  (let ((out (make-tensor (shape tensor) :dtype (dtype tensor) :order (order tensor) :device 'CUDATensor)))
    (with-lowered-views ((ov out) (tv tensor))
      (ew-binary +copy+ ov tv))
    out))
