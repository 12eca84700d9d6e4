ssh test@127.0.0.1 "kill -9 2957"
kill -9 2946
