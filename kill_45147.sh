ssh test@127.0.0.1 "kill -9 9281"
ssh test@127.0.0.1 "kill -9 9282"
ssh test@127.0.0.1 "kill -9 9283"
kill -9 8873
