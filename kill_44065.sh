ssh test@127.0.0.1 "kill -9 2971"
ssh test@127.0.0.1 "kill -9 2972"
kill -9 2946
